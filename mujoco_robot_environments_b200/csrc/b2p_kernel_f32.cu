// fp32 instantiation of the step kernel (throughput mode; BASELINE configs[2]).
#include "b2p_kernel.cuh"
namespace b2p {
cudaError_t configure_f32(int device, int n_env, LaunchGeom* out) { return step_configure<float>(device, n_env, out); }
cudaError_t launch_f32(const LaunchGeom& lg, const void* d_model, const Globals<float>& g, const StepArgs<float>& a, int* next_env, cudaStream_t st) {
  return step_launch<float>(lg, d_model, g, a, next_env, st);
}
cudaError_t prof_read_f32(unsigned long long* out, int reset) { return step_prof_read(out, reset); }
}  // namespace b2p
