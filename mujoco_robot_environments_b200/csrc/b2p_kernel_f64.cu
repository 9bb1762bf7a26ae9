// fp64 instantiation of the step kernel (parity mode; BASELINE configs[1]).
#include "b2p_kernel.cuh"
namespace b2p {
cudaError_t configure_f64(int device, int n_env, LaunchGeom* out) { return step_configure<double>(device, n_env, out); }
cudaError_t launch_f64(const LaunchGeom& lg, const void* d_model, const Globals<double>& g, const StepArgs<double>& a, int* next_env, cudaStream_t st) {
  return step_launch<double>(lg, d_model, g, a, next_env, st);
}
cudaError_t prof_read_f64(unsigned long long* out, int reset) { return step_prof_read(out, reset); }
}  // namespace b2p
