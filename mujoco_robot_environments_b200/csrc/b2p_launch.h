// Entry points of the two kernel translation units (b2p_kernel_f32.cu / b2p_kernel_f64.cu), used by b2p_api.cu.
#pragma once
#include <cuda_runtime.h>

#include "b2p_scratch.h"

namespace b2p {
struct LaunchGeom { int grid, threads, smem_bytes, blocks_per_sm, warps_per_cta; };
cudaError_t configure_f32(int device, int n_env, LaunchGeom* out);
cudaError_t configure_f64(int device, int n_env, LaunchGeom* out);
cudaError_t prof_read_f32(unsigned long long* out, int reset);
cudaError_t prof_read_f64(unsigned long long* out, int reset);
cudaError_t launch_f32(const LaunchGeom& lg, const void* d_model, const Globals<float>& g, const StepArgs<float>& a, int* next_env, cudaStream_t st);
cudaError_t launch_f64(const LaunchGeom& lg, const void* d_model, const Globals<double>& g, const StepArgs<double>& a, int* next_env, cudaStream_t st);
}  // namespace b2p
