// TEST INFRASTRUCTURE ONLY.  Compiles the kernel's per-env templates (csrc/b2p_core.h) for the host with a
// one-thread "block" so the CPU test-suite can check kernel logic against the oracle without a GPU.
// The package never loads this library; the product path is the CUDA build and has no CPU fallback.
#include <cmath>
#include <cstdlib>
#include <new>
using std::sqrt; using std::fabs; using std::sin; using std::cos; using std::pow;
#include "../../mujoco_robot_environments_b200/csrc/b2p_core.h"

using namespace b2p;

template <typename T>
struct Emu {
  DevModel<T> model; HostTables ht;
  std::vector<T> mesh_vert; std::vector<short> g1, g2; std::vector<PairParam<T>> pp; std::vector<NbrPt<T>> nbr_pt; std::vector<NbrSpan> nbr_span; std::vector<T> jovf, geom_dop, xpool, w6; std::vector<GeomRaw<T>> geom_raw;
  Globals<T> g; Scratch<T, false>* s; Scratch<T, true>* sc;   // box-only and class-capable instantiations of the stepping code
};

template <typename T>
static Emu<T>* make(const void* blob, size_t n) {
  DevModel<double> md; Emu<T>* e = new Emu<T>();
  std::string err = build_model(blob, n, md, e->ht);
  if (!err.empty()) { fprintf(stderr, "hostemu: %s\n", err.c_str()); delete e; return nullptr; }
  convert_model(md, e->model);
  e->mesh_vert.assign(e->ht.mesh_vert.begin(), e->ht.mesh_vert.end());
  for (size_t i = 0; i < e->ht.pair_g1.size(); i++) {
    e->g1.push_back((short)e->ht.pair_g1[i]); e->g2.push_back((short)e->ht.pair_g2[i]);
    PairParam<T> p; const double* q = &e->ht.pair_param[12 * i];
    for (int k = 0; k < 3; k++) p.friction[k] = (T)q[k];
    for (int k = 0; k < 2; k++) p.solref[k] = (T)q[3 + k];
    for (int k = 0; k < 5; k++) p.solimp[k] = (T)q[5 + k];
    p.margin = (T)q[10]; p.gap = (T)q[11]; e->pp.push_back(p);
  }
  e->g.mesh_vert = e->mesh_vert.data(); e->g.nbradr = e->ht.mesh_nbradr.data(); e->g.nbrnum = e->ht.mesh_nbrnum.data(); e->g.nbr = e->ht.mesh_nbr.data(); e->g.geom_start = e->ht.geom_start.data();
  e->geom_dop.assign(e->ht.geom_dop.begin(), e->ht.geom_dop.end()); e->g.geom_dop = e->geom_dop.data();
  e->nbr_pt.resize(e->ht.mesh_nbr.size()); e->nbr_span.resize(e->ht.mesh_nbradr.size());
  for (size_t q = 0; q < e->nbr_pt.size(); q++) { const int u = e->ht.mesh_nbr[q]; e->nbr_pt[q].x = (T)e->ht.mesh_vert[3 * u]; e->nbr_pt[q].y = (T)e->ht.mesh_vert[3 * u + 1]; e->nbr_pt[q].z = (T)e->ht.mesh_vert[3 * u + 2]; e->nbr_pt[q].idx = u; }
  for (size_t v = 0; v < e->nbr_span.size(); v++) { e->nbr_span[v].adr = e->ht.mesh_nbradr[v]; e->nbr_span[v].num = e->ht.mesh_nbrnum[v]; }
  e->g.nbr_pt = e->nbr_pt.data(); e->g.nbr_span = e->nbr_span.data();
  e->g.pair_g1 = e->g1.data(); e->g.pair_g2 = e->g2.data(); e->g.pair_param = e->pp.data();
  e->jovf.assign(Pool<T>::JP + 16, T(0)); e->g.jovf = e->jovf.data();
  e->xpool.assign(Pool<T>::JP, T(0)); e->g.xpool = e->xpool.data();
  e->geom_raw.resize(e->ht.geom_raw.size());
  for (size_t i = 0; i < e->geom_raw.size(); i++) {
    const GeomRaw<double>& a = e->ht.geom_raw[i]; GeomRaw<T>& b = e->geom_raw[i];
    b.priority = a.priority; b.condim = a.condim;
    for (int k = 0; k < 3; k++) b.friction[k] = (T)a.friction[k];
    for (int k = 0; k < 2; k++) b.solref[k] = (T)a.solref[k];
    for (int k = 0; k < 5; k++) b.solimp[k] = (T)a.solimp[k];
    b.solmix = (T)a.solmix; b.margin = (T)a.margin; b.gap = (T)a.gap;
  }
  e->g.geom_raw = e->geom_raw.data();
  e->w6.assign((size_t)MAXCON * W6_STRIDE, T(0)); e->g.w6pool = e->w6.data();
  e->s = new Scratch<T, false>(); e->s->jovf = e->jovf.data(); e->s->xpool = e->xpool.data(); e->s->w6 = e->w6.data();
  e->sc = new Scratch<T, true>(); e->sc->jovf = e->jovf.data(); e->sc->xpool = e->xpool.data(); e->sc->w6 = e->w6.data();
  return e;
}

struct EmuArgs {   // mirrors the ctypes struct in tests/hostemu.py
  int mode, n_env, n_steps, control_steps, state_stride, full_contacts, settle_min_steps, out_stride;
  double settle_qvel_tol, settle_qacc_tol;
  void *state, *ctrl, *target, *grip, *prop_size, *prop_active, *env_mask, *converged, *status, *steps_done, *out, *gains;
  int solver; void* prop_type;
};

template <typename T>
static int run(Emu<T>* e, const EmuArgs* a) {
  StepArgs<T> s; std::memset(&s, 0, sizeof s);
  s.mode = a->mode; s.n_env = a->n_env; s.n_steps = a->n_steps; s.control_steps = a->control_steps; s.state_stride = a->state_stride; s.full_contacts = a->full_contacts;
  s.state = (T*)a->state; s.ctrl = (const T*)a->ctrl; s.target = (const T*)a->target; s.grip = (const unsigned char*)a->grip;
  s.prop_size = (const T*)a->prop_size; s.prop_active = (const unsigned char*)a->prop_active; s.env_mask = (const unsigned char*)a->env_mask;
  s.converged = (unsigned char*)a->converged; s.status = (int*)a->status; s.steps_done = (int*)a->steps_done;
  s.settle_qvel_tol = (T)a->settle_qvel_tol; s.settle_qacc_tol = (T)a->settle_qacc_tol; s.settle_min_steps = a->settle_min_steps;
  s.out = (T*)a->out; s.out_stride = a->out_stride; s.out_mask = OUT_ALL; s.gains = (const T*)a->gains; s.prop_type = (const unsigned char*)a->prop_type;
  Ctx c = {0, 1};
  for (int env = 0; env < a->n_env; env++) {
    if (s.env_mask && !s.env_mask[env]) continue;
    if (a->prop_type || model_needs_classes(e->model)) run_env<T, VAR_STEP>(e->model, e->g, *e->sc, s, env, c, false);      // as step_launch chooses
    else if (a->solver == 1) run_env<T, VAR_PGS>(e->model, e->g, *e->s, s, env, c, false);
    else run_env<T, VAR_STEP>(e->model, e->g, *e->s, s, env, c, false);
  }
  return 0;
}

extern "C" {
void* emu_create(const void* blob, size_t n, int dtype) { return dtype == 0 ? (void*)make<double>(blob, n) : (void*)make<float>(blob, n); }
int emu_run(void* h, int dtype, const EmuArgs* a) { return dtype == 0 ? run((Emu<double>*)h, a) : run((Emu<float>*)h, a); }
void emu_osc_config(void* h, int dtype, const double* gains6, const double* null7, double pt, double ot) {
  if (dtype == 0) { auto* e = (Emu<double>*)h; for (int i = 0; i < 6; i++) e->model.osc_gains[i] = gains6[i]; for (int i = 0; i < 7; i++) e->model.osc_null[i] = null7[i]; e->model.osc_pos_thresh = pt; e->model.osc_ori_thresh = ot; }
  else { auto* e = (Emu<float>*)h; for (int i = 0; i < 6; i++) e->model.osc_gains[i] = (float)gains6[i]; for (int i = 0; i < 7; i++) e->model.osc_null[i] = (float)null7[i]; e->model.osc_pos_thresh = (float)pt; e->model.osc_ori_thresh = (float)ot; }
}
int emu_out_size() { return OutLayout::SIZE; }
int emu_scratch_bytes(int dtype) { return dtype == 0 ? (int)sizeof(Scratch<double>) : (int)sizeof(Scratch<float>); }
int emu_model_bytes(int dtype) { return dtype == 0 ? (int)sizeof(DevModel<double>) : (int)sizeof(DevModel<float>); }
void emu_destroy(void* h, int dtype) { if (dtype == 0) { auto* e = (Emu<double>*)h; delete e->s; delete e->sc; delete e; } else { auto* e = (Emu<float>*)h; delete e->s; delete e->sc; delete e; } }
}
// narrow-phase census of a run (tools/narrow_census.py; built with -DB2P_HOST_COUNTERS): support calls, hill-climb steps, MPR queries
extern "C" void emu_newton(long* out) {
#ifdef B2P_HOST_COUNTERS
  out[0] = b2p::g_newton_iters; out[1] = b2p::g_newton_solves; b2p::g_newton_iters = b2p::g_newton_solves = 0;
#else
  out[0] = out[1] = -1;
#endif
}
extern "C" void emu_pair_mpr(long* out) {
#ifdef B2P_HOST_COUNTERS
  for (int i = 0; i < 1024; i++) { out[i] = b2p::g_pair_mpr[i]; b2p::g_pair_mpr[i] = 0; }
#else
  for (int i = 0; i < 1024; i++) out[i] = 0;
#endif
}
extern "C" void emu_counters(long* out, int reset) {
#ifdef B2P_HOST_COUNTERS
  out[0] = b2p::g_support_calls; out[1] = b2p::g_climb_steps; out[2] = b2p::g_mpr_calls; out[3] = b2p::g_cand; out[4] = b2p::g_bb_calls; out[5] = b2p::g_steps;
  if (reset) b2p::g_support_calls = b2p::g_climb_steps = b2p::g_mpr_calls = b2p::g_cand = b2p::g_bb_calls = b2p::g_steps = 0;
#else
  for (int i = 0; i < 6; i++) out[i] = -1; (void)reset;
#endif
}
