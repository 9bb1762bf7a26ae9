// libb2p.so: the C-ABI of include/b2p.h (handle lifetime, uploads, launches).  The step kernel itself lives in
// b2p_kernel.cuh and is instantiated per dtype in b2p_kernel_f32.cu / b2p_kernel_f64.cu.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/b2p.h"
#include "b2p_launch.h"

using namespace b2p;

static_assert((int)OUT_SITE == (int)B2P_OUT_SITE && (int)OUT_CONTACT == (int)B2P_OUT_CONTACT && (int)OUT_MARM == (int)B2P_OUT_MARM && (int)OUT_GEOM == (int)B2P_OUT_GEOM &&
              (int)OUT_BIAS == (int)B2P_OUT_BIAS && (int)OUT_DYN == (int)B2P_OUT_DYN && (int)OUT_EFC == (int)B2P_OUT_EFC && (int)OUT_ALL == (int)B2P_OUT_ALL, "include/b2p.h and b2p_scratch.h disagree on the output groups");
static thread_local std::string g_err;
static int fail(int code, const std::string& msg) { g_err = msg; return code; }
namespace b2p { int api_fail(int code, const char* msg) { return fail(code, msg); } }   // shared with b2p_render.cu
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return fail(B2P_ERR_CUDA, std::string(#x) + ": " + cudaGetErrorString(e_)); } while (0)

// register-resident FMA chain for the FP pipe roofline denominator
template <typename T>
__global__ void fma_peak_kernel(T* out, int iters) {
  T a0 = T(threadIdx.x) * T(1e-3), a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  const T b = T(0.999), cc = T(1e-4);
  for (int i = 0; i < iters; i++) {
    a0 = a0 * b + cc; a1 = a1 * b + cc; a2 = a2 * b + cc; a3 = a3 * b + cc; a4 = a4 * b + cc; a5 = a5 * b + cc; a6 = a6 * b + cc; a7 = a7 * b + cc;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

// ------------------------------------------------------------------------------------------ scene queue order
// The CTA's warps walk the stages of a step in lockstep, so a round costs what its slowest scene costs, and the launch ends
// when the last round does.  Scenes are therefore handed out heaviest first, neighbours in the queue having done similar work
// in the previous launch (work proxy written by the kernel): a counting sort over 256 buckets, one block.  The order within a
// bucket is arbitrary; results do not depend on which warp steps a scene.
__global__ void __launch_bounds__(1024) b2p_order_kernel(const int* __restrict__ work, int n, int* __restrict__ order) {
  __shared__ int hist[256]; __shared__ float total;
  for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0;
  if (threadIdx.x == 0) total = 0.f;
  __syncthreads();
  float part = 0.f;
  for (int e = threadIdx.x; e < n; e += blockDim.x) part += (float)max(work[e], 0);
  for (int o = 16; o; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
  if ((threadIdx.x & 31) == 0) atomicAdd(&total, part);
  __syncthreads();
  const float scale = 96.f * (float)n / fmaxf(total, 1.f);      // bucket width = mean work / 96; everything above 2.6 x the mean shares the top bucket
  auto bucket = [&](int e) { return min(255, (int)((float)max(work[e], 0) * scale)); };
  for (int e = threadIdx.x; e < n; e += blockDim.x) atomicAdd(&hist[bucket(e)], 1);
  __syncthreads();
  if (threadIdx.x == 0) { int acc = 0; for (int b = 255; b >= 0; b--) { const int k = hist[b]; hist[b] = acc; acc += k; } }
  __syncthreads();
  for (int e = threadIdx.x; e < n; e += blockDim.x) order[atomicAdd(&hist[bucket(e)], 1)] = e;
}

// ------------------------------------------------------------------------------------------ scratch of the no-allocation step (b2p_step_physics_noalloc)
// The step kernel's per-warp scratch in global memory is the Jacobian overflow area (rows beyond the shared-memory pool) and the
// scene-queue counter.  Handles get theirs from cudaMalloc in b2p_create; the no-allocation entry point uses THIS array instead:
// one per GPU that loads the library, held until exit, shared by every call (calls that overlap on the GPU race).
// Size: one overflow area per resident warp, at most B2P_NOALLOC_MAX_ENV scenes rounded up to whole CTAs, in the wider dtype.
constexpr size_t NOALLOC_WARPS = B2P_NOALLOC_MAX_ENV + 12;                                    // ceil(n / warps per CTA) CTAs of <= 12 warps
constexpr size_t NOALLOC_AREA = sizeof(double) * Pool<double>::JP > sizeof(float) * Pool<float>::JP ? sizeof(double) * Pool<double>::JP : sizeof(float) * Pool<float>::JP;
constexpr size_t NOALLOC_JOVF_OFF = 256;                                                       // [0, 4): queue counter; the areas start on the next 256-byte line
constexpr size_t NOALLOC_BYTES = NOALLOC_JOVF_OFF + NOALLOC_WARPS * NOALLOC_AREA + 16 * sizeof(double);   // + 16 reals: row_dot's guarded slots
static_assert(NOALLOC_BYTES <= (4u << 20), "the static scratch is meant to stay modest (4 MiB)");
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];

// In host code `g_noalloc_scratch` is only a stand-in for the symbol: its device address has to be asked for, once per device, with
// that device current (b2p_create does, after its cudaSetDevice).
constexpr int NOALLOC_MAXDEV = 64;
static std::mutex g_noalloc_mutex;
static void* g_noalloc_ptr[NOALLOC_MAXDEV];
static int noalloc_resolve(int device) {
  if (device < 0 || device >= NOALLOC_MAXDEV) return B2P_OK;             // such a device simply has no no-allocation step
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (g_noalloc_ptr[device]) return B2P_OK;
  void* p = nullptr;
  cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
  if (e != cudaSuccess || !p) return fail(B2P_ERR_CUDA, std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e));
  g_noalloc_ptr[device] = p;
  return B2P_OK;
}

// ------------------------------------------------------------------------------------------ handle
struct b2p_handle_s {
  int dtype, device, n_env, full_contacts, nq, nv, nu, nprop, ngeom;
  LaunchGeom lg;
  long long launches, order_launches;
  DevModel<double> hmodel;
  void* d_model; void* d_mesh_vert; int *d_nbradr, *d_nbrnum, *d_nbr, *d_geom_start; void* d_geom_dop; void* d_geom_raw; const unsigned char* prop_type; void *d_nbr_pt, *d_nbr_span; short *d_g1, *d_g2; void* d_pair_param;
  void* state; void* out; void* own_out; const void* prop_size; const unsigned char* prop_active; int* d_status;
  void* d_zero_ctrl; int* d_next_env; void* d_jovf; void* d_xpool; void* d_w6pool; int solver;
  const void* env_gains; const unsigned char* env_mask;
  int *d_work, *d_order; int work_steps, sort_queue, out_mask;   // work_steps: physics steps behind d_work (0 = none yet)
};

template <typename T>
static int upload(b2p_handle_s* h, const HostTables& ht) {
  DevModel<T> dm; convert_model(h->hmodel, dm);
  CK(cudaMalloc(&h->d_model, sizeof dm)); CK(cudaMemcpy(h->d_model, &dm, sizeof dm, cudaMemcpyHostToDevice));
  std::vector<T> mv(ht.mesh_vert.begin(), ht.mesh_vert.end());
  CK(cudaMalloc(&h->d_mesh_vert, mv.size() * sizeof(T) + 16)); CK(cudaMemcpy(h->d_mesh_vert, mv.data(), mv.size() * sizeof(T), cudaMemcpyHostToDevice));
  std::vector<PairParam<T>> pp(ht.pair_g1.size()); std::vector<short> g1(pp.size()), g2(pp.size());
  for (size_t i = 0; i < pp.size(); i++) {
    const double* q = &ht.pair_param[12 * i];
    for (int k = 0; k < 3; k++) pp[i].friction[k] = (T)q[k];
    for (int k = 0; k < 2; k++) pp[i].solref[k] = (T)q[3 + k];
    for (int k = 0; k < 5; k++) pp[i].solimp[k] = (T)q[5 + k];
    pp[i].margin = (T)q[10]; pp[i].gap = (T)q[11]; g1[i] = (short)ht.pair_g1[i]; g2[i] = (short)ht.pair_g2[i];
  }
  CK(cudaMalloc(&h->d_pair_param, pp.size() * sizeof(PairParam<T>) + 16)); CK(cudaMemcpy(h->d_pair_param, pp.data(), pp.size() * sizeof(PairParam<T>), cudaMemcpyHostToDevice));
  CK(cudaMalloc(&h->d_g1, g1.size() * 2 + 16)); CK(cudaMemcpy(h->d_g1, g1.data(), g1.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&h->d_g2, g2.size() * 2 + 16)); CK(cudaMemcpy(h->d_g2, g2.data(), g2.size() * 2, cudaMemcpyHostToDevice));
  { std::vector<T> dop(ht.geom_dop.begin(), ht.geom_dop.end()); CK(cudaMalloc(&h->d_geom_dop, dop.size() * sizeof(T) + 16)); CK(cudaMemcpy(h->d_geom_dop, dop.data(), dop.size() * sizeof(T), cudaMemcpyHostToDevice)); }
  {
    std::vector<GeomRaw<T>> gr(ht.geom_raw.size());
    for (size_t i = 0; i < gr.size(); i++) {
      const GeomRaw<double>& a = ht.geom_raw[i]; GeomRaw<T>& b = gr[i];
      b.priority = a.priority; b.condim = a.condim;
      for (int k = 0; k < 3; k++) b.friction[k] = (T)a.friction[k];
      for (int k = 0; k < 2; k++) b.solref[k] = (T)a.solref[k];
      for (int k = 0; k < 5; k++) b.solimp[k] = (T)a.solimp[k];
      b.solmix = (T)a.solmix; b.margin = (T)a.margin; b.gap = (T)a.gap;
    }
    CK(cudaMalloc(&h->d_geom_raw, gr.size() * sizeof(GeomRaw<T>) + 16)); CK(cudaMemcpy(h->d_geom_raw, gr.data(), gr.size() * sizeof(GeomRaw<T>), cudaMemcpyHostToDevice));
  }
  CK(cudaMalloc(&h->d_geom_start, ht.geom_start.size() * 4 + 16)); CK(cudaMemcpy(h->d_geom_start, ht.geom_start.data(), ht.geom_start.size() * 4, cudaMemcpyHostToDevice));
  {
    std::vector<NbrPt<T>> np(ht.mesh_nbr.size()); std::vector<NbrSpan> ns(ht.mesh_nbradr.size());
    for (size_t q = 0; q < np.size(); q++) { const int u = ht.mesh_nbr[q]; np[q].x = (T)ht.mesh_vert[3 * u]; np[q].y = (T)ht.mesh_vert[3 * u + 1]; np[q].z = (T)ht.mesh_vert[3 * u + 2]; np[q].idx = u; }
    for (size_t v = 0; v < ns.size(); v++) { ns[v].adr = ht.mesh_nbradr[v]; ns[v].num = ht.mesh_nbrnum[v]; }
    CK(cudaMalloc(&h->d_nbr_pt, np.size() * sizeof(NbrPt<T>) + 32)); CK(cudaMemcpy(h->d_nbr_pt, np.data(), np.size() * sizeof(NbrPt<T>), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&h->d_nbr_span, ns.size() * sizeof(NbrSpan) + 16)); CK(cudaMemcpy(h->d_nbr_span, ns.data(), ns.size() * sizeof(NbrSpan), cudaMemcpyHostToDevice));
  }
  CK(cudaMalloc(&h->d_nbradr, ht.mesh_nbradr.size() * 4 + 16)); CK(cudaMemcpy(h->d_nbradr, ht.mesh_nbradr.data(), ht.mesh_nbradr.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&h->d_nbrnum, ht.mesh_nbrnum.size() * 4 + 16)); CK(cudaMemcpy(h->d_nbrnum, ht.mesh_nbrnum.data(), ht.mesh_nbrnum.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&h->d_nbr, ht.mesh_nbr.size() * 4 + 16)); CK(cudaMemcpy(h->d_nbr, ht.mesh_nbr.data(), ht.mesh_nbr.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&h->own_out, (size_t)h->n_env * OutLayout::SIZE * sizeof(T))); CK(cudaMemset(h->own_out, 0, (size_t)h->n_env * OutLayout::SIZE * sizeof(T)));
  CK(cudaMalloc(&h->d_zero_ctrl, (size_t)h->n_env * MAXU * sizeof(T))); CK(cudaMemset(h->d_zero_ctrl, 0, (size_t)h->n_env * MAXU * sizeof(T)));
  h->out = h->own_out;
  CK(cudaMalloc(&h->d_next_env, sizeof(int)));
  CK(sizeof(T) == 8 ? configure_f64(h->device, h->n_env, &h->lg) : configure_f32(h->device, h->n_env, &h->lg));
  CK(cudaMalloc(&h->d_jovf, ((size_t)h->lg.grid * h->lg.warps_per_cta * Pool<T>::JP + 16) * sizeof(T)));   // + 16: row_dot reads a tree row in 16 guarded slots   // Jacobian overflow: one area per resident warp
  CK(cudaMalloc(&h->d_w6pool, (size_t)h->lg.grid * h->lg.warps_per_cta * MAXCON * W6_STRIDE * sizeof(T)));
  CK(cudaMalloc(&h->d_work, (size_t)h->n_env * 4)); CK(cudaMemset(h->d_work, 0, (size_t)h->n_env * 4));
  CK(cudaMalloc(&h->d_order, (size_t)h->n_env * 4));
  { const char* sq = getenv("B2P_SORT_QUEUE"); h->sort_queue = sq ? atoi(sq) : (h->n_env > h->lg.grid * h->lg.warps_per_cta); }   // pays off once there is more than one round
  return B2P_OK;
}

template <typename T>
static int launch(b2p_handle_s* h, StepArgs<T>& a, cudaStream_t st) {
  if (!h->state) return fail(B2P_ERR_STATE, "b2p_bind_state has not been called");
  a.n_env = h->n_env; a.state = (T*)h->state; a.state_stride = REC_SIZE; a.prop_size = (const T*)h->prop_size; a.prop_active = h->prop_active; a.prop_type = h->prop_type;
  a.status = h->d_status; a.out = (T*)h->out; a.out_stride = OutLayout::SIZE; a.out_mask = h->out_mask; a.gains = (const T*)h->env_gains;
  Globals<T> g = {(const T*)h->d_mesh_vert, h->d_nbradr, h->d_nbrnum, h->d_nbr, h->d_g1, h->d_g2, (const PairParam<T>*)h->d_pair_param, h->d_geom_start, (const NbrPt<T>*)h->d_nbr_pt, (const NbrSpan*)h->d_nbr_span, (T*)h->d_jovf, (T*)h->d_w6pool, (T*)h->d_xpool, (const GeomRaw<T>*)h->d_geom_raw, (const T*)h->d_geom_dop};
  a.work = h->d_work; a.order = nullptr; a.solver = h->solver; a.need_classes = model_needs_classes(h->hmodel);
  if (h->sort_queue && h->work_steps > 0) {
    b2p_order_kernel<<<1, 1024, 0, st>>>(h->d_work, h->n_env, h->d_order);
    CK(cudaGetLastError());
    a.order = h->d_order; h->order_launches++;
  }
  if ((a.mode == MODE_PHYSICS || a.mode == MODE_OSC) && a.n_steps > 0) h->work_steps = a.mode == MODE_OSC ? a.n_steps * a.control_steps : a.n_steps;
  if constexpr (sizeof(T) == 8) CK(launch_f64(h->lg, h->d_model, g, a, h->d_next_env, st));
  else CK(launch_f32(h->lg, h->d_model, g, a, h->d_next_env, st));
  h->launches++;
  return B2P_OK;
}

template <typename T>
static int run_mode(b2p_handle h, int mode, int n_steps, int control_steps, int full, const void* ctrl, const void* target, const uint8_t* grip,
                    const uint8_t* mask, uint8_t* conv, int32_t* steps_done, double vt, double at, int min_steps, void* stream) {
  StepArgs<T> a; std::memset(&a, 0, sizeof a);
  { const char* ls = getenv("B2P_LOCKSTEP"); a.lockstep = ls ? atoi(ls) : LK_DEFAULT; }
  { const char* lg = getenv("B2P_LOCKGROUP"); a.lock_group = lg ? atoi(lg) : 0; }
  a.mode = mode; a.n_steps = n_steps; a.control_steps = control_steps; a.full_contacts = full;
  a.ctrl = (const T*)(ctrl ? ctrl : h->d_zero_ctrl); a.target = (const T*)target; a.grip = grip; a.env_mask = mask; a.converged = conv; a.steps_done = steps_done;
  a.settle_qvel_tol = (T)vt; a.settle_qacc_tol = (T)at; a.settle_min_steps = min_steps;
  return launch<T>(h, a, (cudaStream_t)stream);
}
template <typename T>
static int place_mode(b2p_handle h, const int32_t* prop, const void* box, const uint8_t* mask, int max_attempts, double dist, int table_geom, unsigned seed, void* out, void* stream, int yaw = 0, void* quat_out = nullptr) {
  StepArgs<T> a; std::memset(&a, 0, sizeof a);
  { const char* ls = getenv("B2P_LOCKSTEP"); a.lockstep = ls ? atoi(ls) : LK_DEFAULT; }
  a.mode = MODE_PLACE; a.n_steps = max_attempts; a.control_steps = 1; a.full_contacts = 1; a.env_mask = mask;
  a.place_prop = prop; a.place_box = (const T*)box; a.place_dist = (T)dist; a.place_table = table_geom; a.place_seed = seed; a.place_out = (T*)out; a.place_yaw = yaw; a.place_quat = (T*)quat_out;
  return launch<T>(h, a, (cudaStream_t)stream);
}
// the step launch of b2p_step_physics with the kernel's scratch taken from g_noalloc_scratch: no queue ordering (its work figures
// belong to the handle's own launches), no handle field is modified
template <typename T>
static int launch_noalloc(b2p_handle_s* h, void* scratch, void* state, const void* ctrl, int n_steps, cudaStream_t st) {
  StepArgs<T> a; std::memset(&a, 0, sizeof a);
  { const char* ls = getenv("B2P_LOCKSTEP"); a.lockstep = ls ? atoi(ls) : LK_DEFAULT; }
  { const char* lg = getenv("B2P_LOCKGROUP"); a.lock_group = lg ? atoi(lg) : 0; }
  a.mode = MODE_PHYSICS; a.n_steps = n_steps; a.control_steps = 1; a.full_contacts = h->full_contacts;
  a.ctrl = (const T*)(ctrl ? ctrl : h->d_zero_ctrl); a.env_mask = h->env_mask;
  a.n_env = h->n_env; a.state = (T*)state; a.state_stride = REC_SIZE; a.prop_size = (const T*)h->prop_size; a.prop_active = h->prop_active; a.prop_type = nullptr;
  a.status = h->d_status; a.out = (T*)h->out; a.out_stride = OutLayout::SIZE; a.out_mask = h->out_mask; a.gains = (const T*)h->env_gains;
  a.work = nullptr; a.order = nullptr; a.solver = B2P_SOLVER_NEWTON; a.need_classes = model_needs_classes(h->hmodel);
  int* next_env = (int*)scratch;                                                     // cleared by the launch itself (cudaMemsetAsync on the stream)
  T* jovf = (T*)((unsigned char*)scratch + NOALLOC_JOVF_OFF);                         // no need to clear: a row is written before it is read
  Globals<T> g = {(const T*)h->d_mesh_vert, h->d_nbradr, h->d_nbrnum, h->d_nbr, h->d_g1, h->d_g2, (const PairParam<T>*)h->d_pair_param, h->d_geom_start, (const NbrPt<T>*)h->d_nbr_pt, (const NbrSpan*)h->d_nbr_span, jovf, nullptr, nullptr, (const GeomRaw<T>*)h->d_geom_raw, (const T*)h->d_geom_dop};
  if constexpr (sizeof(T) == 8) CK(launch_f64(h->lg, h->d_model, g, a, next_env, st));
  else CK(launch_f32(h->lg, h->d_model, g, a, next_env, st));
  return B2P_OK;
}

extern "C" {

const char* b2p_last_error(void) { return g_err.c_str(); }

int b2p_create(const void* blob, size_t nbytes, int n_env, int dtype, int device, b2p_handle* out) {
  if (!blob || !out || n_env < 1 || (dtype != B2P_F64 && dtype != B2P_F32)) return fail(B2P_ERR_ARG, "b2p_create: bad argument");
  b2p_handle_s* h = new b2p_handle_s(); std::memset((void*)h, 0, sizeof *h);
  HostTables ht;
  std::string err = build_model(blob, nbytes, h->hmodel, ht);
  if (!err.empty()) { delete h; return fail(B2P_ERR_MODEL, "b2p_create: " + err); }
  h->out_mask = OUT_ALL;
  h->dtype = dtype; h->device = device; h->n_env = n_env; h->nq = h->hmodel.nq; h->nv = h->hmodel.nv; h->nu = h->hmodel.nu; h->nprop = h->hmodel.nprop; h->ngeom = h->hmodel.ngeom;
  CK(cudaSetDevice(device));
  { const int rc0 = noalloc_resolve(device); if (rc0 != B2P_OK) { delete h; return rc0; } }
  CK(cudaMalloc(&h->d_status, (size_t)n_env * 4)); CK(cudaMemset(h->d_status, 0, (size_t)n_env * 4));
  int rc = dtype == B2P_F64 ? upload<double>(h, ht) : upload<float>(h, ht);
  if (rc != B2P_OK) { delete h; return rc; }
  *out = h;
  return B2P_OK;
}
int b2p_destroy(b2p_handle h) {
  if (!h) return B2P_OK;
  cudaSetDevice(h->device);
  void* ptrs[] = {h->d_model, h->d_mesh_vert, h->d_geom_start, h->d_geom_dop, h->d_geom_raw, h->d_nbr_pt, h->d_nbr_span, h->d_nbradr, h->d_nbrnum, h->d_nbr, h->d_g1, h->d_g2, h->d_pair_param, h->own_out, h->d_status, h->d_zero_ctrl, h->d_next_env, h->d_work, h->d_order, h->d_jovf, h->d_xpool, h->d_w6pool};
  for (void* p : ptrs) if (p) cudaFree(p);
  delete h;
  return B2P_OK;
}
int b2p_query(b2p_handle h, int what) {
  if (!h) return B2P_ERR_ARG;
  switch (what) {
    case B2P_Q_NQ: return h->nq; case B2P_Q_NV: return h->nv; case B2P_Q_NU: return h->nu; case B2P_Q_NPROP: return h->nprop;
    case B2P_Q_STATE_STRIDE: return REC_SIZE; case B2P_Q_OUT_STRIDE: return OutLayout::SIZE; case B2P_Q_MAXPROP: return MAXP;
    case B2P_Q_REC_QPOS: return REC_QPOS; case B2P_Q_REC_QVEL: return REC_QVEL; case B2P_Q_REC_QACC_WARMSTART: return REC_QACC; case B2P_Q_REC_TIME: return REC_TIME;
    case B2P_Q_LAUNCHES: return (int)h->launches; case B2P_Q_SMEM_BYTES: return h->lg.smem_bytes; case B2P_Q_BLOCKS_PER_SM: return h->lg.blocks_per_sm;
    case B2P_Q_THREADS: return h->lg.threads; case B2P_Q_GRID: return h->lg.grid; case B2P_Q_MAXCON: return MAXCON; case B2P_Q_NGEOM: return h->ngeom;
    case B2P_Q_ORDER_LAUNCHES: return (int)h->order_launches;
  }
  return B2P_ERR_ARG;
}
int b2p_out_offset(int field, int* offset, int* count) {
  typedef OutLayout L;
  static const int off[B2P_O_NFIELD] = {L::SITE_XPOS, L::SITE_XMAT, L::QACC, L::QFRC_BIAS, L::CTRL, L::NCON, L::NEFC, L::NITER, L::CONTACT, L::MARM, L::GPOS, L::GMAT, L::QACC_SMOOTH, L::EFC_FORCE};
  static const int cnt[B2P_O_NFIELD] = {6, 18, MAXNV, MAXNV, MAXU, 1, 1, 1, MAXCON * L::CONTACT_W, MAXTD * MAXTD, MAXG * 3, MAXG * 9, MAXNV, MAXEFC_OUT};
  if (field < 0 || field >= B2P_O_NFIELD) return fail(B2P_ERR_ARG, "b2p_out_offset: unknown field");
  if (offset) *offset = off[field];
  if (count) *count = cnt[field];
  return B2P_OK;
}
int b2p_bind_state(b2p_handle h, void* state) {
  if (!h || !state) return fail(B2P_ERR_ARG, "b2p_bind_state: null");
  if (((uintptr_t)state) % 16) return fail(B2P_ERR_ARG, "b2p_bind_state: state must be 16-byte aligned (TMA)");
  h->state = state; return B2P_OK;
}
int b2p_bind_out(b2p_handle h, void* out) { if (!h) return B2P_ERR_ARG; h->out = out ? out : h->own_out; return B2P_OK; }
int b2p_set_out_mask(b2p_handle h, int mask) { if (!h || mask < 0 || mask > OUT_ALL) return fail(B2P_ERR_ARG, "b2p_set_out_mask: bad argument"); h->out_mask = mask; return B2P_OK; }
int b2p_set_solver(b2p_handle h, int solver) {
  if (!h || (solver != B2P_SOLVER_NEWTON && solver != B2P_SOLVER_PGS)) return fail(B2P_ERR_ARG, "b2p_set_solver: bad argument");
  if (solver == B2P_SOLVER_PGS && model_needs_classes(h->hmodel)) return fail(B2P_ERR_ARG, "b2p_set_solver: the PGS instantiation covers the elliptic cone and box / hull geoms only");
  if (solver == B2P_SOLVER_PGS && h->prop_type) return fail(B2P_ERR_ARG, "b2p_set_solver: PGS covers condim 1 and 3; it is not offered while prop classes are bound (b2p_set_env_prop_types)");
  if (solver == B2P_SOLVER_PGS && !h->d_xpool) {   // M^-1 J' per resident warp, J's packed layout
    CK(cudaSetDevice(h->device));
    const size_t w = h->dtype == B2P_F64 ? 8 : 4, jp = h->dtype == B2P_F64 ? Pool<double>::JP : Pool<float>::JP;
    CK(cudaMalloc(&h->d_xpool, (size_t)h->lg.grid * h->lg.warps_per_cta * jp * w));
  }
  h->solver = solver; return B2P_OK;
}
int b2p_set_env_params(b2p_handle h, const void* prop_size, const uint8_t* prop_active) { if (!h) return B2P_ERR_ARG; h->prop_size = prop_size; h->prop_active = prop_active; return B2P_OK; }
int b2p_set_env_prop_types(b2p_handle h, const uint8_t* prop_type) {
  if (!h) return B2P_ERR_ARG;
  if (prop_type && h->hmodel.cone == B2P_CONE_PYRAMIDAL) return fail(B2P_ERR_ARG, "b2p_set_env_prop_types: prop classes (condim-6 contacts) are not offered with a pyramidal cone");
  if (prop_type && h->solver == B2P_SOLVER_PGS) return fail(B2P_ERR_ARG, "b2p_set_env_prop_types: not offered with the PGS solver (condim 1 and 3 only)");
  h->prop_type = prop_type; return B2P_OK;
}
int b2p_set_env_gains(b2p_handle h, const void* gains) { if (!h) return B2P_ERR_ARG; h->env_gains = gains; return B2P_OK; }
int b2p_set_env_mask(b2p_handle h, const uint8_t* mask) { if (!h) return B2P_ERR_ARG; h->env_mask = mask; return B2P_OK; }
int b2p_set_full_contacts(b2p_handle h, int full) { if (!h) return B2P_ERR_ARG; h->full_contacts = full; return B2P_OK; }
int b2p_osc_config(b2p_handle h, const double g[6], const double q0[7], double pt, double ot) {
  if (!h || !g || !q0) return fail(B2P_ERR_ARG, "b2p_osc_config: null");
  for (int i = 0; i < 6; i++) h->hmodel.osc_gains[i] = g[i];
  for (int i = 0; i < 7; i++) h->hmodel.osc_null[i] = q0[i];
  h->hmodel.osc_pos_thresh = pt; h->hmodel.osc_ori_thresh = ot;
  CK(cudaSetDevice(h->device));
  if (h->dtype == B2P_F64) { DevModel<double> dm; convert_model(h->hmodel, dm); CK(cudaMemcpy(h->d_model, &dm, sizeof dm, cudaMemcpyHostToDevice)); }
  else { DevModel<float> dm; convert_model(h->hmodel, dm); CK(cudaMemcpy(h->d_model, &dm, sizeof dm, cudaMemcpyHostToDevice)); }
  return B2P_OK;
}

#define DISPATCH(...) (h->dtype == B2P_F64 ? run_mode<double>(__VA_ARGS__) : run_mode<float>(__VA_ARGS__))

int b2p_step_physics(b2p_handle h, const void* ctrl, int n_steps, void* stream) {
  if (!h || n_steps < 0) return fail(B2P_ERR_ARG, "b2p_step_physics: bad argument");
  return DISPATCH(h, MODE_PHYSICS, n_steps, 1, h->full_contacts, ctrl, nullptr, nullptr, h->env_mask, nullptr, nullptr, 0, 0, 0, stream);
}
int b2p_step_physics_noalloc(b2p_handle h, void* state, const void* ctrl, int n_steps, void* stream) {
  if (!h || !state || n_steps < 0) return fail(B2P_ERR_ARG, "b2p_step_physics_noalloc: bad argument");
  if (((uintptr_t)state) % 16) return fail(B2P_ERR_ARG, "b2p_step_physics_noalloc: state must be 16-byte aligned (TMA)");
  if (h->n_env > B2P_NOALLOC_MAX_ENV)
    return fail(B2P_ERR_ARG, "b2p_step_physics_noalloc: the handle holds " + std::to_string(h->n_env) + " scenes, the static scratch covers at most " + std::to_string((int)B2P_NOALLOC_MAX_ENV) + " (B2P_NOALLOC_MAX_ENV); use b2p_step_physics");
  if ((size_t)h->lg.grid * h->lg.warps_per_cta > NOALLOC_WARPS) return fail(B2P_ERR_ARG, "b2p_step_physics_noalloc: more resident warps than the static scratch has areas for");
  if (h->solver != B2P_SOLVER_NEWTON) return fail(B2P_ERR_ARG, "b2p_step_physics_noalloc: Newton only (the PGS solver's pool is allocated on demand)");
  if (h->prop_type) return fail(B2P_ERR_ARG, "b2p_step_physics_noalloc: box props only (no prop classes bound)");
  void* scratch = h->device >= 0 && h->device < NOALLOC_MAXDEV ? g_noalloc_ptr[h->device] : nullptr;      // resolved by b2p_create on this device
  if (!scratch) return fail(B2P_ERR_STATE, "b2p_step_physics_noalloc: the scratch address was not resolved for this device");
  return h->dtype == B2P_F64 ? launch_noalloc<double>(h, scratch, state, ctrl, n_steps, (cudaStream_t)stream)
                             : launch_noalloc<float>(h, scratch, state, ctrl, n_steps, (cudaStream_t)stream);
}
int b2p_step_osc(b2p_handle h, const void* target13, const uint8_t* grip, int n_ticks, int control_steps, uint8_t* converged, void* stream) {
  if (!h || !target13 || n_ticks < 0 || control_steps < 1) return fail(B2P_ERR_ARG, "b2p_step_osc: bad argument");
  return DISPATCH(h, MODE_OSC, n_ticks, control_steps, h->full_contacts, nullptr, target13, grip, h->env_mask, converged, nullptr, 0, 0, 0, stream);
}
int b2p_place_search(b2p_handle h, const int32_t* prop, const void* box, const uint8_t* mask, int max_attempts, double dist, int table_geom, unsigned seed, void* out, void* stream) {
  if (!h || !prop || !box || !out || max_attempts < 0) return fail(B2P_ERR_ARG, "b2p_place_search: bad argument");
  return h->dtype == B2P_F64 ? place_mode<double>(h, prop, box, mask, max_attempts, dist, table_geom, seed, out, stream)
                             : place_mode<float>(h, prop, box, mask, max_attempts, dist, table_geom, seed, out, stream);
}
int b2p_place_props(b2p_handle h, const int32_t* prop, const void* box, const uint8_t* mask, int max_attempts, double dist, int table_geom, unsigned seed, void* out, void* quat_out, void* stream) {
  if (!h || !prop || !box || !out || !quat_out || max_attempts < 0) return fail(B2P_ERR_ARG, "b2p_place_props: bad argument");
  return h->dtype == B2P_F64 ? place_mode<double>(h, prop, box, mask, max_attempts, dist, table_geom, seed, out, stream, 1, quat_out)
                             : place_mode<float>(h, prop, box, mask, max_attempts, dist, table_geom, seed, out, stream, 1, quat_out);
}
int b2p_settle(b2p_handle h, const uint8_t* mask, int max_steps, int min_steps, double vt, double at, int32_t* steps_done, void* stream) {
  if (!h || max_steps < 0) return fail(B2P_ERR_ARG, "b2p_settle: bad argument");
  return DISPATCH(h, MODE_SETTLE, max_steps, 1, h->full_contacts, nullptr, nullptr, nullptr, mask, nullptr, steps_done, vt, at, min_steps, stream);
}
int b2p_forward(b2p_handle h, int full_contacts, void* stream) {
  if (!h) return fail(B2P_ERR_ARG, "b2p_forward: null handle");
  return DISPATCH(h, MODE_FORWARD, 0, 1, full_contacts, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, 0, 0, stream);
}
int b2p_get(b2p_handle h, int field, void* dst, void* stream) {
  int off, cnt;
  if (!h || !dst) return fail(B2P_ERR_ARG, "b2p_get: null");
  if (b2p_out_offset(field, &off, &cnt) != B2P_OK) return B2P_ERR_ARG;
  const size_t w = h->dtype == B2P_F64 ? 8 : 4;
  CK(cudaMemcpy2DAsync(dst, cnt * w, (const char*)h->out + off * w, OutLayout::SIZE * w, cnt * w, h->n_env, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return B2P_OK;
}
int b2p_status(b2p_handle h, int32_t* flags, int clear, void* stream) {
  if (!h) return B2P_ERR_ARG;
  if (flags) CK(cudaMemcpyAsync(flags, h->d_status, (size_t)h->n_env * 4, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  if (clear) CK(cudaMemsetAsync(h->d_status, 0, (size_t)h->n_env * 4, (cudaStream_t)stream));
  return B2P_OK;
}
int b2p_status_set(b2p_handle h, const int32_t* flags, void* stream) {
  if (!h || !flags) return fail(B2P_ERR_ARG, "b2p_status_set: null");
  CK(cudaMemcpyAsync(h->d_status, flags, (size_t)h->n_env * 4, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return B2P_OK;
}
int b2p_prof_read(b2p_handle h, unsigned long long* out, int reset) {
  if (!h || !out) return fail(B2P_ERR_ARG, "b2p_prof_read: null");
  CK(cudaSetDevice(h->device)); CK(cudaDeviceSynchronize());
  CK(h->dtype == B2P_F64 ? prof_read_f64(out, reset) : prof_read_f32(out, reset));
  return B2P_OK;
}
int b2p_fma_peak(int dtype, int device, double* tflops) {
  if (!tflops) return B2P_ERR_ARG;
  CK(cudaSetDevice(device));
  int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  const int blocks = sms * 8, threads = 256, iters = 1 << 16;
  void* buf; CK(cudaMalloc(&buf, (size_t)blocks * threads * 8));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 4; rep++) {
    CK(cudaEventRecord(e0));
    if (dtype == B2P_F64) fma_peak_kernel<double><<<blocks, threads>>>((double*)buf, iters); else fma_peak_kernel<float><<<blocks, threads>>>((float*)buf, iters);
    CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (rep > 0 && ms < best) best = ms;
  }
  *tflops = 2.0 * 8 * (double)iters * blocks * threads / (best * 1e-3) / 1e12;
  cudaFree(buf); cudaEventDestroy(e0); cudaEventDestroy(e1);
  return B2P_OK;
}
}  // extern "C"
