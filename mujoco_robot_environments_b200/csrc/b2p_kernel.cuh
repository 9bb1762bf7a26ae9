// The sm_100a step kernel (one template, instantiated for float in b2p_kernel_f32.cu and for double in
// b2p_kernel_f64.cu so the two long compiles run in parallel).
//
// Launch shape: persistent grid, one WARP per scene ("warp-per-env").  A CTA holds kWarps warps; model constants
// (DevModel<T>) are staged into shared memory ONCE per CTA with a 1-D TMA bulk copy (cp.async.bulk -> UBLKCP)
// completing on an mbarrier and are shared by the CTA's warps.  Each warp owns one Scratch<T> in shared memory,
// pulls scene indices from a global atomic counter (dynamic load balance: per-scene cost varies with contacts and
// Newton iterations), TMA-loads the scene's 128-real state record into the head of its Scratch and TMA-stores it
// back when the requested steps are done.  HBM therefore sees exactly one coalesced record read + write per scene
// per launch, however many physics steps the launch runs.  Warps never synchronise with each other after staging.
#pragma once
#include <cuda_runtime.h>

#include <cstddef>

#include "b2p_core.h"
#include "b2p_launch.h"

namespace b2p {



#ifndef B2P_WARPS_F32
#define B2P_WARPS_F32 12
#endif
template <typename T> struct KernelCfg { static constexpr int kWarps = B2P_WARPS_F32; static constexpr int kMinBlocks = 1; };
#ifndef B2P_WARPS_F64
#define B2P_WARPS_F64 6
#endif
template <> struct KernelCfg<double> { static constexpr int kWarps = B2P_WARPS_F64; static constexpr int kMinBlocks = 1; };

// ------------------------------------------------------------------------------------------ TMA / mbarrier helpers
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned phase) {
  unsigned done;
  do {
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(smem_u32(bar)), "r"(phase) : "memory");
  } while (!done);
}
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_store_1d(void* dst, const void* src, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

template <typename T, bool C = false>
struct alignas(16) SmemLayout {
  DevModel<T> model;
  alignas(16) Scratch<T, C> scratch[KernelCfg<T>::kWarps];
  alignas(8) unsigned long long bar[1 + KernelCfg<T>::kWarps];
};

static_assert(sizeof(SmemLayout<float>) + 128 <= 232448 && sizeof(SmemLayout<double>) + 128 <= 232448 && sizeof(SmemLayout<float, true>) == sizeof(SmemLayout<float, false>), "shared memory per CTA exceeds the 227 KB an sm_100 CTA may use: shrink Pool<T>::JPS or the warps per CTA");

template <typename T, int kVar, bool C>
__global__ void __launch_bounds__(32 * KernelCfg<T>::kWarps, KernelCfg<T>::kMinBlocks)
b2p_step_kernel(const DevModel<T>* __restrict__ gmodel, Globals<T> g, StepArgs<T> a, int* __restrict__ next_env) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  SmemLayout<T, C>& sm = *reinterpret_cast<SmemLayout<T, C>*>(smem_raw);
  constexpr unsigned kModelBytes = (unsigned)sizeof(DevModel<T>);
  constexpr unsigned kRecBytes = REC_SIZE * sizeof(T);
  static_assert(sizeof(DevModel<T>) % 16 == 0, "DevModel must be a multiple of 16 bytes for cp.async.bulk");
  typedef Scratch<T, C> ScratchT;
  static_assert(offsetof(ScratchT, qpos) == 0 && sizeof(ScratchT) % 16 == 0, "state record must head a 16-byte aligned Scratch");
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // lockstep groups: the CTA's warps walk the stage barriers in groups of a.lock_group warps (named barriers 1..), so that a
  // warp waits for the slowest solve of its group only; the whole CTA meets once per scene at the queue
  const int lg = (a.lock_group > 0 && a.lock_group < KernelCfg<T>::kWarps) ? a.lock_group : 0;
  int grp_n = 0, grp_id = 0;
  if (lg) { grp_id = warp / lg; const int w0 = grp_id * lg, w1 = (w0 + lg < KernelCfg<T>::kWarps) ? w0 + lg : KernelCfg<T>::kWarps; grp_n = 32 * (w1 - w0); grp_id += 1; }
  const Ctx c = {lane, 32, a.lockstep, grp_id, grp_n};
  if (threadIdx.x == 0) {
    for (int i = 0; i <= KernelCfg<T>::kWarps; i++) mbar_init(&sm.bar[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) { mbar_expect_tx(&sm.bar[0], kModelBytes); tma_load_1d(&sm.model, gmodel, kModelBytes, &sm.bar[0]); }
  mbar_wait(&sm.bar[0], 0);
  Scratch<T, C>& s = sm.scratch[warp];
  constexpr bool kPlace = kVar == VAR_PLACE;
  if (lane == 0) { s.jovf = g.jovf + (size_t)(blockIdx.x * KernelCfg<T>::kWarps + warp) * Pool<T>::JP; s.xpool = g.xpool ? g.xpool + (size_t)(blockIdx.x * KernelCfg<T>::kWarps + warp) * Pool<T>::JP : nullptr; s.w6 = g.w6pool + (size_t)(blockIdx.x * KernelCfg<T>::kWarps + warp) * MAXCON * W6_STRIDE; }
  __syncwarp();
  unsigned long long* bar = &sm.bar[1 + warp];
  unsigned phase = 0;
  for (;;) {   // one round = one scene per warp; in lockstep mode the CTA's warps walk the round's stages together
    int env = 0;
    if (lane == 0) {
      do { env = atomicAdd(next_env, 1); if (a.order && env < a.n_env) env = a.order[env]; } while (env < a.n_env && a.env_mask && !a.env_mask[env]);
    }
    env = __shfl_sync(0xffffffffu, env, 0);
    const bool active = env < a.n_env;
    if (c.lock ? !__syncthreads_or(active) : !active) break;
    T* rec = a.state + (size_t)(active ? env : 0) * a.state_stride;
    if (active) {
      if (lane == 0) { fence_async_smem(); mbar_expect_tx(bar, kRecBytes); tma_load_1d(s.qpos, rec, kRecBytes, bar); }
      mbar_wait(bar, phase); phase ^= 1;
    }
    PROF_T0
    run_env<T, kVar>(sm.model, g, s, a, env, c, true, active);
    if (active) PROF_ADD(PF_TOTAL)
    if (active) {
      __syncwarp();
      if (!kPlace && lane == 0) { fence_async_smem(); tma_store_1d(rec, s.qpos, kRecBytes); }
      __syncwarp();
    }
  }
}

// stage cycle counters of this translation unit (all zero unless built with -DB2P_PROF)
static cudaError_t step_prof_read(unsigned long long* out, int reset) {
#ifdef B2P_PROF
  cudaError_t e = cudaMemcpyFromSymbol(out, g_prof, sizeof(unsigned long long) * PF_N);
  if (e != cudaSuccess) return e;
  if (reset) { unsigned long long z[PF_N] = {0}; e = cudaMemcpyToSymbol(g_prof, z, sizeof z); }
  return e;
#else
  (void)reset; for (int i = 0; i < PF_N; i++) out[i] = 0;
  return cudaSuccess;
#endif
}

template <typename T> static size_t step_smem_size() { return sizeof(SmemLayout<T, false>) + 128; }

// occupancy query + attribute setup (once per handle)
template <typename T>
static cudaError_t step_configure(int device, int n_env, LaunchGeom* out) {
  const size_t smem = step_smem_size<T>();
  cudaError_t e = cudaFuncSetAttribute(b2p_step_kernel<T, VAR_STEP, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(b2p_step_kernel<T, VAR_PLACE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(b2p_step_kernel<T, VAR_PGS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(b2p_step_kernel<T, VAR_STEP, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(b2p_step_kernel<T, VAR_PLACE, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  int per_sm = 0, sms = 0;
  const int threads = 32 * KernelCfg<T>::kWarps;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, b2p_step_kernel<T, VAR_STEP, false>, threads, smem);
  if (e != cudaSuccess) return e;
  e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  if (e != cudaSuccess) return e;
  if (per_sm < 1) return cudaErrorLaunchOutOfResources;
  out->threads = threads; out->smem_bytes = (int)smem; out->blocks_per_sm = per_sm; out->warps_per_cta = KernelCfg<T>::kWarps;
  out->grid = sms * per_sm;
  const int need = (n_env + KernelCfg<T>::kWarps - 1) / KernelCfg<T>::kWarps;
  if (out->grid > need) out->grid = need;
  return cudaSuccess;
}

template <typename T>
static cudaError_t step_launch(const LaunchGeom& lg, const void* d_model, const Globals<T>& g, const StepArgs<T>& a, int* next_env, cudaStream_t st) {
  cudaError_t e = cudaMemsetAsync(next_env, 0, sizeof(int), st);
  if (e != cudaSuccess) return e;
  // five instantiations: {step, place} x {box props only, prop classes} + PGS (box props only).  The box-only step kernel is the
  // throughput path; a handle with prop classes bound (b2p_set_env_prop_types) launches the class-capable ones.
  const bool cls = a.prop_type != nullptr || a.need_classes != 0;      // pyramidal-cone rows and tree cylinders live in the class-capable instantiations only
  const DevModel<T>* dm = (const DevModel<T>*)d_model;
  if (a.mode == MODE_PLACE) { if (cls) b2p_step_kernel<T, VAR_PLACE, true><<<lg.grid, lg.threads, lg.smem_bytes, st>>>(dm, g, a, next_env); else b2p_step_kernel<T, VAR_PLACE, false><<<lg.grid, lg.threads, lg.smem_bytes, st>>>(dm, g, a, next_env); }
  else if (a.solver == 1 && !cls) b2p_step_kernel<T, VAR_PGS, false><<<lg.grid, lg.threads, lg.smem_bytes, st>>>(dm, g, a, next_env);
  else if (cls) b2p_step_kernel<T, VAR_STEP, true><<<lg.grid, lg.threads, lg.smem_bytes, st>>>(dm, g, a, next_env);
  else b2p_step_kernel<T, VAR_STEP, false><<<lg.grid, lg.threads, lg.smem_bytes, st>>>(dm, g, a, next_env);
  return cudaGetLastError();
}

}  // namespace b2p
