/* TEST INFRASTRUCTURE -- NOT A PRODUCT PATH.  See b2p_oracle.h.
 *
 * Plain C99, fp64, one scene at a time, clarity over speed.  Each stage names the MuJoCo
 * 3.2.7 routine it restates (third-party, un-vendored: reference pyproject.toml:42; call
 * sites mujoco_robot_environments/models/robot_arm.py:78-79) and the SURVEY.md Appendix B
 * item that summarises its semantics.  PARITY UNPINNED (no reference golden vectors).
 */
#include "b2p_oracle.h"
#include "../include/b2p_model_layout.h"
#include "../include/b2p_prop_types.h"
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define MINVAL 1e-15
#define MAXNV 64
#define MAXBODY 64
#define MAXGEOM 64
#define MAXSITE 32
#define MAXPROP 8
#define MINIMP 0.0001
#define MAXIMP 0.9999

typedef struct {
  double dist, pos[3], frame[9], includemargin, friction[5], solref[2], solimp[5], mu;
  int dim, geom1, geom2, exclude, efc_address;
} contact_t;

enum { ROW_EQ, ROW_LIMIT, ROW_CONTACT };
enum { ST_SATISFIED, ST_QUADRATIC, ST_CONE };

struct orc_t {
  unsigned char* blob; unsigned char* blob0;   /* blob: this instance's copy (prop geoms re-classed by orc_set_prop_type); blob0: as compiled */
  const int* hdr; const double* opt; const int* table;
  int nq, nv, nu, nbody, njnt, ngeom, nsite, neq, npair, nmesh, nprop;
  double timestep, gravity[3], impratio, tolerance, ls_tolerance, ccd_tolerance, meaninertia;
  int iterations, ls_iterations, integrator, cone, ccd_iterations, full_contacts, solver;
  /* per-scene mutable model params (props) */
  double geom_size[3 * MAXGEOM], geom_rbound[MAXGEOM], body_mass[MAXBODY], body_inertia[3 * MAXBODY];
  double body_invweight0[2 * MAXBODY], geom_invweight0[2 * MAXGEOM], dof_invweight0[MAXNV];
  double prop_size[3 * MAXPROP]; int prop_active[MAXPROP], prop_type[MAXPROP];
  double geom_half[3 * MAXGEOM];   /* bounding half extents in the geom frame (mid phase of the timing build) */
  /* state */
  double qpos[MAXNV + 8], qvel[MAXNV], ctrl[16], qacc_warmstart[MAXNV], time;
  /* derived */
  double qacc[MAXNV], qacc_smooth[MAXNV], qfrc_bias[MAXNV], qfrc_passive[MAXNV], qfrc_actuator[MAXNV];
  double qfrc_smooth[MAXNV], qfrc_constraint[MAXNV], act_force[16], act_length[16], act_velocity[16];
  double xpos[3 * MAXBODY], xquat[4 * MAXBODY], xmat[9 * MAXBODY], xipos[3 * MAXBODY], ximat[9 * MAXBODY];
  double xanchor[3 * MAXNV], xaxis[3 * MAXNV];
  double geom_xpos[3 * MAXGEOM], geom_xmat[9 * MAXGEOM], site_xpos[3 * MAXSITE], site_xmat[9 * MAXSITE];
  double subtree_com[3 * MAXBODY], subtree_mass[MAXBODY], cinert[10 * MAXBODY], crb[10 * MAXBODY];
  double cdof[6 * MAXNV], cdof_dot[6 * MAXNV], cvel[6 * MAXBODY];
  double qM[MAXNV * MAXNV], qLD[MAXNV * MAXNV], qH[MAXNV * MAXNV];
  int dof_active[MAXNV];
  contact_t con[ORC_MAXCON]; int ncon;
  int nefc, efc_type[ORC_MAXEFC], efc_id[ORC_MAXEFC], efc_state[ORC_MAXEFC];
  double* efc_J; /* nefc x nv dense */
  int efc_nnz[ORC_MAXEFC]; unsigned char* efc_ind; /* per row: indices of the non-zero Jacobian entries (ORC_MAXEFC x MAXNV) */
  double *jacbuf, *solbuf; /* scratch of make_constraint (4 x 3 x nv) and of the solver, allocated once */
  double efc_pos[ORC_MAXEFC], efc_margin[ORC_MAXEFC], efc_D[ORC_MAXEFC], efc_R[ORC_MAXEFC], efc_aref[ORC_MAXEFC];
  double efc_vel[ORC_MAXEFC], efc_force[ORC_MAXEFC], efc_diagApprox[ORC_MAXEFC], efc_KBIP[4 * ORC_MAXEFC];
  double solver_stats[4]; int solver_iter;
  /* OSC */
  double osc_gains[6], osc_null[7], osc_pos_thresh, osc_ori_thresh, osc_tau[7];
};

#define FI(o, f) ((const int*)((o)->blob + (o)->table[2 * (f)]))
#define FD(o, f) ((const double*)((o)->blob + (o)->table[2 * (f)]))

/* ------------------------------------------------------------------ small math (mju_*) */
static double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static void cross3(double* r, const double* a, const double* b) {
  double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
static double norm3(const double* a) { return sqrt(dot3(a, a)); }
static double normalize3(double* a) {
  double n = norm3(a);
  if (n < MINVAL) { a[0] = 1; a[1] = 0; a[2] = 0; return 0; }
  a[0] /= n; a[1] /= n; a[2] /= n; return n;
}
static void mulquat(double* r, const double* a, const double* b) {
  double t[4] = {a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3], a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
                 a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1], a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]};
  memcpy(r, t, sizeof t);
}
static void normquat(double* q) {
  double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < MINVAL) { q[0] = 1; q[1] = q[2] = q[3] = 0; } else { q[0] /= n; q[1] /= n; q[2] /= n; q[3] /= n; }
}
static void quat2mat(double* m, const double* q) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  m[0] = w * w + x * x - y * y - z * z; m[1] = 2 * (x * y - w * z); m[2] = 2 * (x * z + w * y);
  m[3] = 2 * (x * y + w * z); m[4] = w * w - x * x + y * y - z * z; m[5] = 2 * (y * z - w * x);
  m[6] = 2 * (x * z - w * y); m[7] = 2 * (y * z + w * x); m[8] = w * w - x * x - y * y + z * z;
}
static void mulmatvec3(double* r, const double* m, const double* v) {
  double x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2], y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2], z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
static void mulmatTvec3(double* r, const double* m, const double* v) {
  double x = m[0] * v[0] + m[3] * v[1] + m[6] * v[2], y = m[1] * v[0] + m[4] * v[1] + m[7] * v[2], z = m[2] * v[0] + m[5] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
static void rotvecquat(double* r, const double* v, const double* q) {
  double m[9]; quat2mat(m, q); mulmatvec3(r, m, v);
}
/* mju_mat2Quat branch order (App. B item 17) */
static void mat2quat(double* q, const double* m) {
  if (m[0] + m[4] + m[8] > 0) {
    q[0] = 0.5 * sqrt(1 + m[0] + m[4] + m[8]); q[1] = 0.25 * (m[7] - m[5]) / q[0];
    q[2] = 0.25 * (m[2] - m[6]) / q[0]; q[3] = 0.25 * (m[3] - m[1]) / q[0];
  } else if (m[0] > m[4] && m[0] > m[8]) {
    q[1] = 0.5 * sqrt(1 + m[0] - m[4] - m[8]); q[0] = 0.25 * (m[7] - m[5]) / q[1];
    q[2] = 0.25 * (m[1] + m[3]) / q[1]; q[3] = 0.25 * (m[2] + m[6]) / q[1];
  } else if (m[4] > m[8]) {
    q[2] = 0.5 * sqrt(1 - m[0] + m[4] - m[8]); q[0] = 0.25 * (m[2] - m[6]) / q[2];
    q[1] = 0.25 * (m[1] + m[3]) / q[2]; q[3] = 0.25 * (m[5] + m[7]) / q[2];
  } else {
    q[3] = 0.5 * sqrt(1 - m[0] - m[4] + m[8]); q[0] = 0.25 * (m[3] - m[1]) / q[3];
    q[1] = 0.25 * (m[2] + m[6]) / q[3]; q[2] = 0.25 * (m[5] + m[7]) / q[3];
  }
  normquat(q);
}
/* mju_makeFrame (App. B item 12): frame[0:3] given, complete y,z deterministically */
static void make_frame(double* f) {
  normalize3(f);
  if (norm3(f + 3) < 0.5) {
    f[3] = f[4] = f[5] = 0;
    if (f[1] < 0.5 && f[1] > -0.5) f[4] = 1; else f[5] = 1;
  }
  double d = dot3(f, f + 3);
  for (int k = 0; k < 3; k++) f[3 + k] -= d * f[k];
  normalize3(f + 3);
  cross3(f + 6, f, f + 3);
}

/* spatial algebra about the subtree-root COM; vectors are [ang; lin] (App. B item 5) */
static void inert_com(double* res, const double* inert, const double* mat, const double* dif, double mass) {
  double tmp[9];
  for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) tmp[3 * r + c] = mat[3 * r + c] * inert[c];
  double I[9];
  for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) I[3 * r + c] = tmp[3 * r] * mat[3 * c] + tmp[3 * r + 1] * mat[3 * c + 1] + tmp[3 * r + 2] * mat[3 * c + 2];
  res[0] = I[0] + mass * (dif[1] * dif[1] + dif[2] * dif[2]); res[1] = I[4] + mass * (dif[0] * dif[0] + dif[2] * dif[2]);
  res[2] = I[8] + mass * (dif[0] * dif[0] + dif[1] * dif[1]); res[3] = I[1] - mass * dif[0] * dif[1];
  res[4] = I[2] - mass * dif[0] * dif[2]; res[5] = I[5] - mass * dif[1] * dif[2];
  res[6] = mass * dif[0]; res[7] = mass * dif[1]; res[8] = mass * dif[2]; res[9] = mass;
}
static void mul_inert_vec(double* r, const double* i, const double* v) {
  r[0] = i[0] * v[0] + i[3] * v[1] + i[4] * v[2] - i[8] * v[4] + i[7] * v[5];
  r[1] = i[3] * v[0] + i[1] * v[1] + i[5] * v[2] + i[8] * v[3] - i[6] * v[5];
  r[2] = i[4] * v[0] + i[5] * v[1] + i[2] * v[2] - i[7] * v[3] + i[6] * v[4];
  r[3] = i[8] * v[1] - i[7] * v[2] + i[9] * v[3];
  r[4] = i[6] * v[2] - i[8] * v[0] + i[9] * v[4];
  r[5] = i[7] * v[0] - i[6] * v[1] + i[9] * v[5];
}
static void cross_motion(double* r, const double* vel, const double* v) {
  double a[3], b[3], c[3];
  cross3(a, vel, v); cross3(b, vel, v + 3); cross3(c, vel + 3, v);
  r[0] = a[0]; r[1] = a[1]; r[2] = a[2]; r[3] = b[0] + c[0]; r[4] = b[1] + c[1]; r[5] = b[2] + c[2];
}
static void cross_force(double* r, const double* vel, const double* f) {
  double a[3], b[3], c[3];
  cross3(a, vel, f); cross3(b, vel + 3, f + 3); cross3(c, vel, f + 3);
  r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2]; r[3] = c[0]; r[4] = c[1]; r[5] = c[2];
}

/* ------------------------------------------------------------------ lifetime */
static void apply_prop_params(orc_t* o) {
  const int* pb = FI(o, B2PM_F_PROP_BODYID); const int* pg = FI(o, B2PM_F_PROP_GEOMID);
  const double* pm = FD(o, B2PM_F_PROP_MASS); const int* pd = FI(o, B2PM_F_PROP_DOFADR);
  for (int p = 0; p < o->nprop; p++) {
    const double* s = o->prop_size + 3 * p; int b = pb[p], g = pg[p]; double m, I[3], half[3];
    b2p_prop_inertia(o->prop_type[p], s, pm[p], &m, I, half);   /* box: m/3 (b^2 + c^2); sphere 2/5 m r^2; cylinder m (r^2/4 + h^2/3), m r^2/2 */
    memcpy(o->geom_size + 3 * g, s, 24); memcpy(o->geom_half + 3 * g, half, 24);
    o->geom_rbound[g] = norm3(half) ; if (o->prop_type[p] == B2P_GEOM_SPHERE) o->geom_rbound[g] = s[0];
    o->body_mass[b] = m;
    memcpy(o->body_inertia + 3 * b, I, 24);
    /* mj_setConst for a lone free body: invweight0 = (1/m, mean(1/I)) (App. B item 9) */
    double tr = 1 / m, rot = (1 / I[0] + 1 / I[1] + 1 / I[2]) / 3;
    o->body_invweight0[2 * b] = tr; o->body_invweight0[2 * b + 1] = rot;
    o->geom_invweight0[2 * g] = tr; o->geom_invweight0[2 * g + 1] = rot;
    for (int k = 0; k < 3; k++) { o->dof_invweight0[pd[p] + k] = tr; o->dof_invweight0[pd[p] + 3 + k] = rot; }
    for (int k = 0; k < 6; k++) o->dof_active[pd[p] + k] = o->prop_active[p];
  }
}
/* re-class prop p's collision geom (box as compiled / sphere / cylinder): the instance's blob copy gets the class's MJCF fields */
void orc_set_prop_type(orc_t* o, int p, int geom_type) {
  if (p < 0 || p >= o->nprop) return;
  const int g = FI(o, B2PM_F_PROP_GEOMID)[p];
  static const int fields[] = {B2PM_F_GEOM_TYPE, B2PM_F_GEOM_CONDIM, B2PM_F_GEOM_PRIORITY, B2PM_F_GEOM_FRICTION, B2PM_F_GEOM_SOLREF, B2PM_F_GEOM_SOLIMP,
                               B2PM_F_GEOM_SOLMIX, B2PM_F_GEOM_MARGIN, B2PM_F_GEOM_GAP};
  static const int width[] = {4, 4, 4, 24, 16, 40, 8, 8, 8};
  for (int f = 0; f < 9; f++) memcpy(o->blob + o->table[2 * fields[f]] + (size_t)width[f] * g, o->blob0 + o->table[2 * fields[f]] + (size_t)width[f] * g, width[f]);
  if (geom_type == B2P_GEOM_SPHERE || geom_type == B2P_GEOM_CYLINDER) {
    const b2p_prop_class* c = geom_type == B2P_GEOM_SPHERE ? &B2P_PROP_CLASS_SPHERE : &B2P_PROP_CLASS_CYLINDER;
    ((int*)(o->blob + o->table[2 * B2PM_F_GEOM_TYPE]))[g] = c->geom_type; ((int*)(o->blob + o->table[2 * B2PM_F_GEOM_CONDIM]))[g] = c->condim;
    ((int*)(o->blob + o->table[2 * B2PM_F_GEOM_PRIORITY]))[g] = c->priority;
    memcpy(o->blob + o->table[2 * B2PM_F_GEOM_FRICTION] + 24 * (size_t)g, c->friction, 24); memcpy(o->blob + o->table[2 * B2PM_F_GEOM_SOLREF] + 16 * (size_t)g, c->solref, 16);
    memcpy(o->blob + o->table[2 * B2PM_F_GEOM_SOLIMP] + 40 * (size_t)g, c->solimp, 40);
    ((double*)(o->blob + o->table[2 * B2PM_F_GEOM_SOLMIX]))[g] = c->solmix; ((double*)(o->blob + o->table[2 * B2PM_F_GEOM_MARGIN]))[g] = c->margin;
    ((double*)(o->blob + o->table[2 * B2PM_F_GEOM_GAP]))[g] = c->gap;
    o->prop_type[p] = geom_type;
  } else o->prop_type[p] = B2P_GEOM_BOX;
  apply_prop_params(o);
}

orc_t* orc_create(const void* blob, size_t nbytes) {
  const int* hdr = (const int*)blob;
  if (nbytes < 1024 || (unsigned)hdr[0] != B2PM_MAGIC || hdr[1] != B2PM_VERSION) return NULL;
  orc_t* o = (orc_t*)calloc(1, sizeof(orc_t));
  o->blob = (unsigned char*)malloc(nbytes); memcpy(o->blob, blob, nbytes);
  o->blob0 = (unsigned char*)malloc(nbytes); memcpy(o->blob0, blob, nbytes);
  o->hdr = (const int*)o->blob; o->opt = (const double*)(o->blob + B2PM_HEADER_BYTES);
  o->table = (const int*)(o->blob + B2PM_TABLE_OFFSET);
  o->nq = hdr[B2PM_C_NQ]; o->nv = hdr[B2PM_C_NV]; o->nu = hdr[B2PM_C_NU]; o->nbody = hdr[B2PM_C_NBODY];
  o->njnt = hdr[B2PM_C_NJNT]; o->ngeom = hdr[B2PM_C_NGEOM]; o->nsite = hdr[B2PM_C_NSITE]; o->neq = hdr[B2PM_C_NEQ];
  o->npair = hdr[B2PM_C_NPAIR]; o->nmesh = hdr[B2PM_C_NMESH]; o->nprop = hdr[B2PM_C_NPROP];
  if (o->nv > MAXNV || o->nbody > MAXBODY || o->ngeom > MAXGEOM || o->nsite > MAXSITE || o->nprop > MAXPROP) { free(o->blob); free(o); return NULL; }
  const double* op = o->opt;
  o->timestep = op[B2PM_O_TIMESTEP]; o->gravity[0] = op[B2PM_O_GRAVITY_X]; o->gravity[1] = op[B2PM_O_GRAVITY_Y]; o->gravity[2] = op[B2PM_O_GRAVITY_Z];
  o->impratio = op[B2PM_O_IMPRATIO]; o->tolerance = op[B2PM_O_TOLERANCE]; o->ls_tolerance = op[B2PM_O_LS_TOLERANCE];
  o->iterations = (int)op[B2PM_O_ITERATIONS]; o->ls_iterations = (int)op[B2PM_O_LS_ITERATIONS];
  o->integrator = (int)op[B2PM_O_INTEGRATOR]; o->cone = (int)op[B2PM_O_CONE];
  o->ccd_iterations = (int)op[B2PM_O_CCD_ITERATIONS]; o->ccd_tolerance = op[B2PM_O_CCD_TOLERANCE]; o->meaninertia = op[B2PM_O_MEANINERTIA];
  o->full_contacts = 1;
  memcpy(o->geom_size, FD(o, B2PM_F_GEOM_SIZE), 24 * o->ngeom);
  memcpy(o->geom_rbound, FD(o, B2PM_F_GEOM_RBOUND), 8 * o->ngeom);
  memcpy(o->body_mass, FD(o, B2PM_F_BODY_MASS), 8 * o->nbody);
  memcpy(o->body_inertia, FD(o, B2PM_F_BODY_INERTIA), 24 * o->nbody);
  memcpy(o->body_invweight0, FD(o, B2PM_F_BODY_INVWEIGHT0), 16 * o->nbody);
  memcpy(o->geom_invweight0, FD(o, B2PM_F_GEOM_INVWEIGHT0), 16 * o->ngeom);
  memcpy(o->dof_invweight0, FD(o, B2PM_F_DOF_INVWEIGHT0), 8 * o->nv);
  for (int i = 0; i < o->nv; i++) o->dof_active[i] = 1;
  const int* pg = FI(o, B2PM_F_PROP_GEOMID);
  { const double* ab = FD(o, B2PM_F_GEOM_AABB); for (int g = 0; g < o->ngeom; g++) memcpy(o->geom_half + 3 * g, ab + 6 * g + 3, 24); }
  for (int p = 0; p < o->nprop; p++) { memcpy(o->prop_size + 3 * p, o->geom_size + 3 * pg[p], 24); o->prop_active[p] = 1; o->prop_type[p] = B2P_GEOM_BOX; }
  apply_prop_params(o);
  o->efc_J = (double*)calloc((size_t)ORC_MAXEFC * MAXNV, sizeof(double));
  o->efc_ind = (unsigned char*)calloc((size_t)ORC_MAXEFC * MAXNV, 1);
  o->jacbuf = (double*)calloc((size_t)12 * MAXNV, sizeof(double));
  o->solbuf = (double*)calloc((size_t)2 * ORC_MAXEFC + 8 * MAXNV + MAXNV * MAXNV, sizeof(double));
  double g[6] = {350, 20, 500, 100, 200, 30}, q0[7] = {0, -0.785, 0, -2.356, 0, 1.571, 0.785};
  orc_osc_config(o, g, q0, 5e-3, 68e-3);
  orc_reset(o);
  return o;
}
void orc_destroy(orc_t* o) { if (o) { free(o->efc_J); free(o->efc_ind); free(o->jacbuf); free(o->solbuf); free(o->blob); free(o->blob0); free(o); } }
void orc_reset(orc_t* o) {
  memcpy(o->qpos, FD(o, B2PM_F_QPOS0), 8 * o->nq);
  memset(o->qvel, 0, sizeof o->qvel); memset(o->ctrl, 0, sizeof o->ctrl);
  memset(o->qacc_warmstart, 0, sizeof o->qacc_warmstart); memset(o->qacc, 0, sizeof o->qacc);
  o->time = 0;
}
void orc_set_prop(orc_t* o, int p, const double half[3], int active) {
  if (p < 0 || p >= o->nprop) return;
  memcpy(o->prop_size + 3 * p, half, 24); o->prop_active[p] = active; apply_prop_params(o);
}
void orc_set_option(orc_t* o, int what, double v) {
  if (what == 0) o->full_contacts = (int)v; else if (what == 1) o->integrator = (int)v;
  else if (what == 2) o->tolerance = v; else if (what == 3) o->iterations = (int)v; else if (what == 4) o->solver = (int)v;
}
int orc_dim(orc_t* o, int w) {
  int d[10] = {o->nq, o->nv, o->nu, o->nbody, o->ngeom, o->nsite, o->ncon, o->nefc, o->nprop, o->solver_iter};
  return (w >= 0 && w < 10) ? d[w] : -1;
}
double* orc_ptr(orc_t* o, int f) {
  switch (f) {
    case ORC_QPOS: return o->qpos; case ORC_QVEL: return o->qvel; case ORC_CTRL: return o->ctrl;
    case ORC_QACC_WARMSTART: return o->qacc_warmstart; case ORC_TIME: return &o->time; case ORC_QACC: return o->qacc;
    case ORC_QACC_SMOOTH: return o->qacc_smooth; case ORC_QFRC_BIAS: return o->qfrc_bias; case ORC_QFRC_PASSIVE: return o->qfrc_passive;
    case ORC_QFRC_ACTUATOR: return o->qfrc_actuator; case ORC_QFRC_SMOOTH: return o->qfrc_smooth; case ORC_QFRC_CONSTRAINT: return o->qfrc_constraint;
    case ORC_XPOS: return o->xpos; case ORC_XQUAT: return o->xquat; case ORC_XMAT: return o->xmat; case ORC_XIPOS: return o->xipos;
    case ORC_GEOM_XPOS: return o->geom_xpos; case ORC_GEOM_XMAT: return o->geom_xmat; case ORC_SITE_XPOS: return o->site_xpos; case ORC_SITE_XMAT: return o->site_xmat;
    case ORC_QM: return o->qM; case ORC_SUBTREE_COM: return o->subtree_com; case ORC_CINERT: return o->cinert; case ORC_CDOF: return o->cdof;
    case ORC_CVEL: return o->cvel; case ORC_CDOF_DOT: return o->cdof_dot;
    case ORC_EFC_J: return o->efc_J; case ORC_EFC_POS: return o->efc_pos; case ORC_EFC_D: return o->efc_D; case ORC_EFC_R: return o->efc_R;
    case ORC_EFC_AREF: return o->efc_aref; case ORC_EFC_FORCE: return o->efc_force; case ORC_EFC_VEL: return o->efc_vel; case ORC_EFC_MARGIN: return o->efc_margin;
    case ORC_PROP_SIZE: return o->prop_size; case ORC_ACT_FORCE: return o->act_force; case ORC_OSC_TAU: return o->osc_tau; case ORC_SOLVER_STATS: return o->solver_stats;
  }
  return NULL;
}

/* ------------------------------------------------------------------ mj_kinematics (App. B item 4) */
static void kinematics(orc_t* o) {
  const int *par = FI(o, B2PM_F_BODY_PARENTID), *jnum = FI(o, B2PM_F_BODY_JNTNUM), *jadr = FI(o, B2PM_F_BODY_JNTADR);
  const int *jtype = FI(o, B2PM_F_JNT_TYPE), *jqadr = FI(o, B2PM_F_JNT_QPOSADR);
  const double *bpos = FD(o, B2PM_F_BODY_POS), *bquat = FD(o, B2PM_F_BODY_QUAT), *bipos = FD(o, B2PM_F_BODY_IPOS), *biquat = FD(o, B2PM_F_BODY_IQUAT);
  const double *jpos = FD(o, B2PM_F_JNT_POS), *jaxis = FD(o, B2PM_F_JNT_AXIS), *qpos0 = FD(o, B2PM_F_QPOS0);
  memset(o->xpos, 0, 24); o->xquat[0] = 1; o->xquat[1] = o->xquat[2] = o->xquat[3] = 0;
  quat2mat(o->xmat, o->xquat); memset(o->xipos, 0, 24); quat2mat(o->ximat, o->xquat);
  for (int i = 1; i < o->nbody; i++) {
    double pos[3], quat[4];
    int p = par[i];
    if (jnum[i] == 1 && jtype[jadr[i]] == B2P_JNT_FREE) {
      int j = jadr[i]; double* q = o->qpos + jqadr[j];
      normquat(q + 3);
      memcpy(pos, q, 24); memcpy(quat, q + 3, 32);
      memcpy(o->xanchor + 3 * j, pos, 24); o->xaxis[3 * j] = 0; o->xaxis[3 * j + 1] = 0; o->xaxis[3 * j + 2] = 1;
    } else {
      mulmatvec3(pos, o->xmat + 9 * p, bpos + 3 * i);
      for (int k = 0; k < 3; k++) pos[k] += o->xpos[3 * p + k];
      mulquat(quat, o->xquat + 4 * p, bquat + 4 * i);
      for (int jj = 0; jj < jnum[i]; jj++) {
        int j = jadr[i] + jj;
        double a[3];
        rotvecquat(a, jpos + 3 * j, quat);
        for (int k = 0; k < 3; k++) o->xanchor[3 * j + k] = pos[k] + a[k];
        rotvecquat(o->xaxis + 3 * j, jaxis + 3 * j, quat);
        double ang = o->qpos[jqadr[j]] - qpos0[jqadr[j]], s = sin(0.5 * ang);
        double qloc[4] = {cos(0.5 * ang), s * jaxis[3 * j], s * jaxis[3 * j + 1], s * jaxis[3 * j + 2]};
        mulquat(quat, quat, qloc);
        rotvecquat(a, jpos + 3 * j, quat);
        for (int k = 0; k < 3; k++) pos[k] = o->xanchor[3 * j + k] - a[k];
      }
    }
    normquat(quat);
    memcpy(o->xpos + 3 * i, pos, 24); memcpy(o->xquat + 4 * i, quat, 32);
    quat2mat(o->xmat + 9 * i, quat);
    double t[3]; mulmatvec3(t, o->xmat + 9 * i, bipos + 3 * i);
    for (int k = 0; k < 3; k++) o->xipos[3 * i + k] = pos[k] + t[k];
    double qi[4]; mulquat(qi, quat, biquat + 4 * i); quat2mat(o->ximat + 9 * i, qi);
  }
  const int* gb = FI(o, B2PM_F_GEOM_BODYID); const double *gp = FD(o, B2PM_F_GEOM_POS), *gq = FD(o, B2PM_F_GEOM_QUAT);
  for (int g = 0; g < o->ngeom; g++) {
    int b = gb[g]; double t[3], q[4];
    mulmatvec3(t, o->xmat + 9 * b, gp + 3 * g);
    for (int k = 0; k < 3; k++) o->geom_xpos[3 * g + k] = o->xpos[3 * b + k] + t[k];
    mulquat(q, o->xquat + 4 * b, gq + 4 * g); quat2mat(o->geom_xmat + 9 * g, q);
  }
  const int* sb = FI(o, B2PM_F_SITE_BODYID); const double *sp = FD(o, B2PM_F_SITE_POS), *sq = FD(o, B2PM_F_SITE_QUAT);
  for (int s = 0; s < o->nsite; s++) {
    int b = sb[s]; double t[3], q[4];
    mulmatvec3(t, o->xmat + 9 * b, sp + 3 * s);
    for (int k = 0; k < 3; k++) o->site_xpos[3 * s + k] = o->xpos[3 * b + k] + t[k];
    mulquat(q, o->xquat + 4 * b, sq + 4 * s); quat2mat(o->site_xmat + 9 * s, q);
  }
}

/* ------------------------------------------------------------------ mj_comPos (App. B item 5) */
static void com_pos(orc_t* o) {
  const int *par = FI(o, B2PM_F_BODY_PARENTID), *root = FI(o, B2PM_F_BODY_ROOTID);
  const int *jnum = FI(o, B2PM_F_BODY_JNTNUM), *jadr = FI(o, B2PM_F_BODY_JNTADR), *jtype = FI(o, B2PM_F_JNT_TYPE), *jdof = FI(o, B2PM_F_JNT_DOFADR);
  int nb = o->nbody;
  for (int i = 0; i < nb; i++) { o->subtree_mass[i] = o->body_mass[i]; for (int k = 0; k < 3; k++) o->subtree_com[3 * i + k] = o->body_mass[i] * o->xipos[3 * i + k]; }
  for (int i = nb - 1; i > 0; i--) { int p = par[i]; o->subtree_mass[p] += o->subtree_mass[i]; for (int k = 0; k < 3; k++) o->subtree_com[3 * p + k] += o->subtree_com[3 * i + k]; }
  for (int i = 0; i < nb; i++) {
    if (o->subtree_mass[i] < MINVAL) memcpy(o->subtree_com + 3 * i, o->xipos + 3 * i, 24);
    else for (int k = 0; k < 3; k++) o->subtree_com[3 * i + k] /= o->subtree_mass[i];
  }
  for (int i = 1; i < nb; i++) {
    double off[3];
    for (int k = 0; k < 3; k++) off[k] = o->xipos[3 * i + k] - o->subtree_com[3 * root[i] + k];
    inert_com(o->cinert + 10 * i, o->body_inertia + 3 * i, o->ximat + 9 * i, off, o->body_mass[i]);
    for (int jj = 0; jj < jnum[i]; jj++) {
      int j = jadr[i] + jj, d = jdof[j];
      for (int k = 0; k < 3; k++) off[k] = o->subtree_com[3 * root[i] + k] - o->xanchor[3 * j + k];
      if (jtype[j] == B2P_JNT_FREE) {
        memset(o->cdof + 6 * d, 0, 18 * 8);
        for (int k = 0; k < 3; k++) o->cdof[6 * (d + k) + 3 + k] = 1;
        for (int k = 0; k < 3; k++) {
          double ax[3] = {o->xmat[9 * i + k], o->xmat[9 * i + 3 + k], o->xmat[9 * i + 6 + k]};
          memcpy(o->cdof + 6 * (d + 3 + k), ax, 24); cross3(o->cdof + 6 * (d + 3 + k) + 3, ax, off);
        }
      } else {
        memcpy(o->cdof + 6 * d, o->xaxis + 3 * j, 24); cross3(o->cdof + 6 * d + 3, o->xaxis + 3 * j, off);
      }
    }
  }
}

/* ------------------------------------------------------------------ mj_crb + mj_factorM (dense storage, tree sparsity) */
static void crb(orc_t* o) {
  const int *par = FI(o, B2PM_F_BODY_PARENTID), *dbody = FI(o, B2PM_F_DOF_BODYID), *dpar = FI(o, B2PM_F_DOF_PARENTID);
  const double* arm = FD(o, B2PM_F_DOF_ARMATURE);
  int nv = o->nv;
  memcpy(o->crb, o->cinert, sizeof(double) * 10 * o->nbody);
  memset(o->crb, 0, 80);
  for (int i = o->nbody - 1; i > 0; i--) if (par[i] > 0) for (int k = 0; k < 10; k++) o->crb[10 * par[i] + k] += o->crb[10 * i + k];
  memset(o->qM, 0, sizeof(double) * nv * nv);
  for (int i = 0; i < nv; i++) {
    double buf[6]; mul_inert_vec(buf, o->crb + 10 * dbody[i], o->cdof + 6 * i);
    for (int j = i; j >= 0; j = dpar[j]) {
      double v = 0; for (int k = 0; k < 6; k++) v += o->cdof[6 * j + k] * buf[k];
      if (j == i) v += arm[i];
      o->qM[i * nv + j] = v; o->qM[j * nv + i] = v;
    }
  }
}
/* L'DL in place on a dense copy: LD[k][i] (i ancestor of k) holds L, diagonal holds D */
static void factor_tree(orc_t* o, double* LD) {
  const int* dpar = FI(o, B2PM_F_DOF_PARENTID); int nv = o->nv;
  for (int k = nv - 1; k >= 0; k--) {
    double dk = LD[k * nv + k];
    for (int i = dpar[k]; i >= 0; i = dpar[i]) {
      double t = LD[k * nv + i] / dk;
      for (int j = i; j >= 0; j = dpar[j]) LD[i * nv + j] -= t * LD[k * nv + j];
      LD[k * nv + i] = t;
    }
  }
}
static void solve_tree(orc_t* o, const double* LD, double* x) {
  const int* dpar = FI(o, B2PM_F_DOF_PARENTID); int nv = o->nv;
  for (int i = nv - 1; i >= 0; i--) for (int j = dpar[i]; j >= 0; j = dpar[j]) x[j] -= LD[i * nv + j] * x[i];
  for (int i = 0; i < nv; i++) x[i] /= LD[i * nv + i];
  for (int i = 0; i < nv; i++) for (int j = dpar[i]; j >= 0; j = dpar[j]) x[i] -= LD[i * nv + j] * x[j];
}

/* ------------------------------------------------------------------ Jacobians (mj_jac) */
static void jac_point(orc_t* o, int body, const double* point, double* jacp, double* jacr) {
  const int *root = FI(o, B2PM_F_BODY_ROOTID), *par = FI(o, B2PM_F_BODY_PARENTID), *dnum = FI(o, B2PM_F_BODY_DOFNUM), *dadr = FI(o, B2PM_F_BODY_DOFADR);
  const int* dpar = FI(o, B2PM_F_DOF_PARENTID);
  int nv = o->nv;
  if (jacp) memset(jacp, 0, 24 * nv);
  if (jacr) memset(jacr, 0, 24 * nv);
  while (body > 0 && dnum[body] == 0) body = par[body];
  if (body <= 0) return;
  double off[3];
  for (int k = 0; k < 3; k++) off[k] = point[k] - o->subtree_com[3 * root[body] + k];
  for (int i = dadr[body] + dnum[body] - 1; i >= 0; i = dpar[i]) {
    const double* c = o->cdof + 6 * i; double t[3];
    if (jacr) for (int k = 0; k < 3; k++) jacr[k * nv + i] = c[k];
    if (jacp) { cross3(t, c, off); for (int k = 0; k < 3; k++) jacp[k * nv + i] = c[3 + k] + t[k]; }
  }
}

/* ------------------------------------------------------------------ collision */
typedef struct { const double* pos; const double* mat; int type; const double* size; const double* vert; int nvert; double margin;
                 const int *nbradr, *nbrnum, *nbr; int cur; } cobj_t;   /* hull graph (mesh-local indices) + warm-start vertex: timing build only */

/* support point of a box / convex mesh in world direction dir (|dir|=1), inflated by margin/2 (App. B item 12).
 * Strict build: exhaustive search over the hull vertices.  Timing build (-DORC_FAST, bench.py's CPU arm): hill climbing over the
 * hull's vertex graph with a warm start inside one query, which is what MuJoCo does for meshes that carry graph data; on a convex
 * hull both reach the same maximum. */
static void support(cobj_t* c, const double* dir, double* res) {
  double ld[3], lp[3];
  mulmatTvec3(ld, c->mat, dir);
  if (c->type == B2P_GEOM_BOX) {
    for (int k = 0; k < 3; k++) lp[k] = ld[k] >= 0 ? c->size[k] : -c->size[k];
  } else if (c->type == B2P_GEOM_SPHERE) {
    for (int k = 0; k < 3; k++) lp[k] = c->size[0] * ld[k];
  } else if (c->type == B2P_GEOM_CYLINDER) {   /* disc of radius size[0] swept along local z by +- size[1] */
    const double rho = sqrt(ld[0] * ld[0] + ld[1] * ld[1]);
    lp[0] = rho > MINVAL ? c->size[0] * ld[0] / rho : 0; lp[1] = rho > MINVAL ? c->size[0] * ld[1] / rho : 0; lp[2] = ld[2] >= 0 ? c->size[1] : -c->size[1];
  } else {
#ifdef ORC_FAST
    int v = c->cur; double bd = dot3(c->vert + 3 * v, ld);
    for (;;) {
      int nb = v;
      for (int q = c->nbradr[v]; q < c->nbradr[v] + c->nbrnum[v]; q++) { int u = c->nbr[q]; double d = dot3(c->vert + 3 * u, ld); if (d > bd) { bd = d; nb = u; } }
      if (nb == v) break;
      v = nb;
    }
    c->cur = v; memcpy(lp, c->vert + 3 * v, 24);
#else
    int best = 0; double bd = -1e300;
    for (int i = 0; i < c->nvert; i++) { double d = dot3(c->vert + 3 * i, ld); if (d > bd) { bd = d; best = i; } }
    memcpy(lp, c->vert + 3 * best, 24);
#endif
  }
  mulmatvec3(res, c->mat, lp);
  for (int k = 0; k < 3; k++) res[k] += c->pos[k] + 0.5 * c->margin * dir[k];
}
typedef struct { double v[3], v1[3], v2[3]; } sup_t;
static void mink_support(cobj_t* a, cobj_t* b, const double* dir, sup_t* s) {
  double nd[3] = {-dir[0], -dir[1], -dir[2]};
  support(a, dir, s->v1); support(b, nd, s->v2);
  for (int k = 0; k < 3; k++) s->v[k] = s->v1[k] - s->v2[k];
}
static int is_zero(double x) { return fabs(x) < 2.220446049250313e-16; }
static void portal_dir(const sup_t* p, double* dir) {
  double a[3], b[3];
  for (int k = 0; k < 3; k++) { a[k] = p[2].v[k] - p[1].v[k]; b[k] = p[3].v[k] - p[1].v[k]; }
  cross3(dir, a, b); normalize3(dir);
}
static int portal_reach_tol(const sup_t* p, const sup_t* v4, const double* dir, double tol) {
  double dv4 = dot3(v4->v, dir), m = dv4 - dot3(p[1].v, dir), d2 = dv4 - dot3(p[2].v, dir), d3 = dv4 - dot3(p[3].v, dir);
  if (d2 < m) m = d2; if (d3 < m) m = d3;
  return is_zero(m - tol) || m < tol;
}
static void expand_portal(sup_t* p, const sup_t* v4) {
  double c[3]; cross3(c, v4->v, p[0].v);
  if (dot3(p[1].v, c) > 0) { if (dot3(p[2].v, c) > 0) p[1] = *v4; else p[3] = *v4; }
  else { if (dot3(p[3].v, c) > 0) p[2] = *v4; else p[1] = *v4; }
}
/* closest point on triangle (a,b,c) to the origin (Ericson) -> squared distance */
static double point_tri_dist2(const double* a, const double* b, const double* c, double* w) {
  double ab[3], ac[3], ap[3], bp[3], cp[3];
  for (int k = 0; k < 3; k++) { ab[k] = b[k] - a[k]; ac[k] = c[k] - a[k]; ap[k] = -a[k]; bp[k] = -b[k]; cp[k] = -c[k]; }
  double d1 = dot3(ab, ap), d2 = dot3(ac, ap), d3 = dot3(ab, bp), d4 = dot3(ac, bp), d5 = dot3(ab, cp), d6 = dot3(ac, cp);
  double v, wgt;
  if (d1 <= 0 && d2 <= 0) { memcpy(w, a, 24); }
  else if (d3 >= 0 && d4 <= d3) { memcpy(w, b, 24); }
  else if (d1 * d4 - d3 * d2 <= 0 && d1 >= 0 && d3 <= 0) { v = d1 / (d1 - d3); for (int k = 0; k < 3; k++) w[k] = a[k] + v * ab[k]; }
  else if (d6 >= 0 && d5 <= d6) { memcpy(w, c, 24); }
  else if (d5 * d2 - d1 * d6 <= 0 && d2 >= 0 && d6 <= 0) { wgt = d2 / (d2 - d6); for (int k = 0; k < 3; k++) w[k] = a[k] + wgt * ac[k]; }
  else if (d3 * d6 - d5 * d4 <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0) { wgt = (d4 - d3) / ((d4 - d3) + (d5 - d6)); for (int k = 0; k < 3; k++) w[k] = b[k] + wgt * (c[k] - b[k]); }
  else {
    double va = d3 * d6 - d5 * d4, vb = d5 * d2 - d1 * d6, vc = d1 * d4 - d3 * d2, den = 1 / (va + vb + vc);
    v = vb * den; wgt = vc * den;
    for (int k = 0; k < 3; k++) w[k] = a[k] + ab[k] * v + ac[k] * wgt;
  }
  return dot3(w, w);
}
/* Minkowski Portal Refinement after libccd's ccdMPRPenetration (what mjc_Convex calls in 3.2.x).
 * returns 1 and (depth, dir [obj1->obj2], pos) if the inflated shapes intersect */
static int mpr_penetration(cobj_t* o1, cobj_t* o2, int maxiter, double tol, double* depth, double* dir, double* pos) {
  sup_t p[4], v4; double d[3], va[3], vb[3];
  /* discoverPortal */
  for (int k = 0; k < 3; k++) { p[0].v1[k] = o1->pos[k]; p[0].v2[k] = o2->pos[k]; p[0].v[k] = p[0].v1[k] - p[0].v2[k]; }
  if (is_zero(p[0].v[0]) && is_zero(p[0].v[1]) && is_zero(p[0].v[2])) p[0].v[0] += 1e-10;
  for (int k = 0; k < 3; k++) d[k] = -p[0].v[k];
  normalize3(d);
  mink_support(o1, o2, d, &p[1]);
  double dt = dot3(p[1].v, d);
  if (is_zero(dt) || dt < 0) return 0;
  cross3(d, p[0].v, p[1].v);
  if (is_zero(dot3(d, d))) {
    if (is_zero(dot3(p[1].v, p[1].v))) { *depth = 0; dir[0] = dir[1] = dir[2] = 0; }       /* touching */
    else { memcpy(dir, p[1].v, 24); *depth = normalize3(dir); }                            /* origin on v0-v1 segment */
    for (int k = 0; k < 3; k++) pos[k] = 0.5 * (p[1].v1[k] + p[1].v2[k]);
    return 1;
  }
  normalize3(d);
  mink_support(o1, o2, d, &p[2]);
  dt = dot3(p[2].v, d);
  if (is_zero(dt) || dt < 0) return 0;
  for (int k = 0; k < 3; k++) { va[k] = p[1].v[k] - p[0].v[k]; vb[k] = p[2].v[k] - p[0].v[k]; }
  cross3(d, va, vb); normalize3(d);
  if (dot3(d, p[0].v) > 0) { sup_t t = p[1]; p[1] = p[2]; p[2] = t; for (int k = 0; k < 3; k++) d[k] = -d[k]; }
  for (int guard = 0;; guard++) {
    if (guard > 100) return 0;
    mink_support(o1, o2, d, &p[3]);
    dt = dot3(p[3].v, d);
    if (is_zero(dt) || dt < 0) return 0;
    int cont = 0;
    cross3(va, p[1].v, p[3].v);
    dt = dot3(va, p[0].v);
    if (dt < 0 && !is_zero(dt)) { p[2] = p[3]; cont = 1; }
    if (!cont) {
      cross3(va, p[3].v, p[2].v);
      dt = dot3(va, p[0].v);
      if (dt < 0 && !is_zero(dt)) { p[1] = p[3]; cont = 1; }
    }
    if (!cont) break;
    for (int k = 0; k < 3; k++) { va[k] = p[1].v[k] - p[0].v[k]; vb[k] = p[2].v[k] - p[0].v[k]; }
    cross3(d, va, vb); normalize3(d);
  }
  /* refinePortal */
  for (int guard = 0;; guard++) {
    portal_dir(p, d);
    dt = dot3(d, p[1].v);
    if (is_zero(dt) || dt > 0) break;  /* portal encapsulates origin */
    mink_support(o1, o2, d, &v4);
    dt = dot3(v4.v, d);
    if (!(is_zero(dt) || dt > 0) || portal_reach_tol(p, &v4, d, tol) || guard > maxiter) return 0;
    expand_portal(p, &v4);
  }
  /* findPenetr */
  for (int it = 0;; it++) {
    portal_dir(p, d);
    mink_support(o1, o2, d, &v4);
    if (portal_reach_tol(p, &v4, d, tol) || it > maxiter) {
      double pd[3];
      *depth = sqrt(point_tri_dist2(p[1].v, p[2].v, p[3].v, pd));
      if (is_zero(pd[0]) && is_zero(pd[1]) && is_zero(pd[2])) memcpy(pd, d, 24);
      normalize3(pd); memcpy(dir, pd, 24);
      /* findPos: barycentric coordinates of the origin in the portal tetrahedron */
      double b[4], c[3], sum;
      cross3(c, p[1].v, p[2].v); b[0] = dot3(c, p[3].v);
      cross3(c, p[3].v, p[2].v); b[1] = dot3(c, p[0].v);
      cross3(c, p[0].v, p[1].v); b[2] = dot3(c, p[3].v);
      cross3(c, p[2].v, p[1].v); b[3] = dot3(c, p[0].v);
      sum = b[0] + b[1] + b[2] + b[3];
      if (is_zero(sum) || sum < 0) {
        b[0] = 0;
        cross3(c, p[2].v, p[3].v); b[1] = dot3(c, d);
        cross3(c, p[3].v, p[1].v); b[2] = dot3(c, d);
        cross3(c, p[1].v, p[2].v); b[3] = dot3(c, d);
        sum = b[1] + b[2] + b[3];
      }
      for (int k = 0; k < 3; k++) {
        double p1 = 0, p2 = 0;
        for (int i = 0; i < 4; i++) { p1 += b[i] * p[i].v1[k]; p2 += b[i] * p[i].v2[k]; }
        pos[k] = 0.5 * (p1 + p2) / sum;
      }
      return 1;
    }
    expand_portal(p, &v4);
  }
}

typedef struct { double dist, pos[3], normal[3]; } rawcon_t;

/* mjc_PlaneBox: up to 4 of the 8 corners within margin */
static int plane_box(const double* ppos, const double* pmat, const double* bpos, const double* bmat, const double* bsize, double margin, rawcon_t* out) {
  double n[3] = {pmat[2], pmat[5], pmat[8]};
  int cnt = 0;
  for (int i = 0; i < 8 && cnt < 4; i++) {
    double loc[3] = {(i & 1 ? bsize[0] : -bsize[0]), (i & 2 ? bsize[1] : -bsize[1]), (i & 4 ? bsize[2] : -bsize[2])}, w[3], dif[3];
    mulmatvec3(w, bmat, loc);
    for (int k = 0; k < 3; k++) { w[k] += bpos[k]; dif[k] = w[k] - ppos[k]; }
    double dist = dot3(dif, n);
    if (dist > margin) continue;
    out[cnt].dist = dist; memcpy(out[cnt].normal, n, 24);
    for (int k = 0; k < 3; k++) out[cnt].pos[k] = w[k] - 0.5 * dist * n[k];
    cnt++;
  }
  return cnt;
}

/* mjc_PlaneConvex (engine_collision_convex.c) for a hull: the support vertex along -normal, then the support vertices of three
 * directions tilted off -normal (120 degrees apart in the plane), kept when they are new vertices inside the margin -- at most 4
 * contacts, all with the plane normal.  RECOLLECTION of MuJoCo 3.2.7: the tilt magnitude is not pinned;
 * B2P_PLANE_TILT is used by the oracle and the kernel alike.  `vid` returns the chosen vertex so that duplicates are dropped by
 * identity, not by distance. */
#define B2P_PLANE_TILT 0.05
static void support_vertex(cobj_t* c, const double* dir, double* res, int* vid) {
  double keep = c->margin; c->margin = 0; support(c, dir, res); c->margin = keep;
  double ld[3]; mulmatTvec3(ld, c->mat, dir);
  int best = 0; double bd = -1e300;
  for (int i = 0; i < c->nvert; i++) { double d = dot3(c->vert + 3 * i, ld); if (d > bd) { bd = d; best = i; } }
  *vid = best;
}
static int plane_convex(const double* ppos, const double* pmat, cobj_t* hull, double margin, rawcon_t* out) {
  double n[3] = {pmat[2], pmat[5], pmat[8]}, t1[3] = {pmat[0], pmat[3], pmat[6]}, t2[3] = {pmat[1], pmat[4], pmat[7]};
  int cnt = 0, seen[4];
  for (int q = 0; q < 4; q++) {
    double dir[3];
    if (q == 0) for (int k = 0; k < 3; k++) dir[k] = -n[k];
    else {
      const double ca = q == 1 ? 1.0 : -0.5, sa = q == 1 ? 0.0 : (q == 2 ? 0.8660254037844387 : -0.8660254037844387);   /* 0, 120, 240 degrees */
      for (int k = 0; k < 3; k++) dir[k] = -n[k] + B2P_PLANE_TILT * (ca * t1[k] + sa * t2[k]);
      normalize3(dir);
    }
    double w[3]; int vid;
    support_vertex(hull, dir, w, &vid);
    int dup = 0;
    for (int i = 0; i < cnt; i++) if (seen[i] == vid) dup = 1;
    if (dup) continue;
    double dif[3] = {w[0] - ppos[0], w[1] - ppos[1], w[2] - ppos[2]};
    const double dist = dot3(dif, n);
    if (dist > margin) continue;
    seen[cnt] = vid;
    out[cnt].dist = dist; memcpy(out[cnt].normal, n, 24);
    for (int k = 0; k < 3; k++) out[cnt].pos[k] = w[k] - 0.5 * dist * n[k];
    cnt++;
  }
  return cnt;
}

/* Box-box: separating-axis test (15 axes) + reference-face clipping (<= 8 points) or edge-edge point.
 * returns #contacts, or -1 when the boxes are separated by 0 < s <= margin (caller falls back to the
 * margin-inflated convex routine, which reports the single closest-feature contact). */
#define BB_TIE 1e-6
static int box_box(const double* p1, const double* m1, const double* s1, const double* p2, const double* m2, const double* s2, double margin, rawcon_t* out) {
  double R[9], aR[9], t[3], dp[3];
  for (int k = 0; k < 3; k++) dp[k] = p2[k] - p1[k];
  mulmatTvec3(t, m1, dp);
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
    R[3 * i + j] = m1[i] * m2[j] + m1[3 + i] * m2[3 + j] + m1[6 + i] * m2[6 + j];
    aR[3 * i + j] = fabs(R[3 * i + j]) + 1e-12;
  }
  double best = -1e300; int code = -1; double bn[3] = {0, 0, 0};
  for (int i = 0; i < 3; i++) { /* faces of box 1 */
    double s = fabs(t[i]) - (s1[i] + s2[0] * aR[3 * i] + s2[1] * aR[3 * i + 1] + s2[2] * aR[3 * i + 2]);
    if (s > margin) return 0;
    if (s > best) { best = s; code = i; }
  }
  for (int j = 0; j < 3; j++) { /* faces of box 2 */
    double tj = t[0] * R[j] + t[1] * R[3 + j] + t[2] * R[6 + j];
    double s = fabs(tj) - (s2[j] + s1[0] * aR[j] + s1[1] * aR[3 + j] + s1[2] * aR[6 + j]);
    if (s > margin) return 0;
    if (s > best + BB_TIE) { best = s; code = 3 + j; }   /* parallel faces tie up to rounding: box 1 keeps the reference face */
  }
  double best_edge = -1e300; int ecode = -1;
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { /* edge i of box1 x edge j of box2, in box-1 coordinates */
    int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    double l = sqrt(R[3 * i1 + j] * R[3 * i1 + j] + R[3 * i2 + j] * R[3 * i2 + j]);
    if (l < 1e-6) continue;
    double tl = t[i2] * R[3 * i1 + j] - t[i1] * R[3 * i2 + j];
    double ra = s1[i1] * aR[3 * i2 + j] + s1[i2] * aR[3 * i1 + j], rb = s2[j1] * aR[3 * i + j2] + s2[j2] * aR[3 * i + j1];
    double s = (fabs(tl) - (ra + rb)) / l;
    if (s > margin) return 0;
    if (s > best_edge) {
      best_edge = s; ecode = 6 + 3 * i + j;
      double a[3] = {m1[i], m1[3 + i], m1[6 + i]}, b[3] = {m2[j], m2[3 + j], m2[6 + j]};
      cross3(bn, a, b); normalize3(bn);
      if (dot3(bn, dp) < 0) for (int k = 0; k < 3; k++) bn[k] = -bn[k];
    }
  }
  if (ecode >= 0 && best_edge > best + 1e-6 * (1 + fabs(best))) { best = best_edge; code = ecode; }
  if (best > 0) return -1;
  if (code >= 6) { /* edge-edge: closest points of the two supporting edges */
    int i = (code - 6) / 3, j = (code - 6) % 3;
    double pa[3], pb[3], ua[3] = {m1[i], m1[3 + i], m1[6 + i]}, ub[3] = {m2[j], m2[3 + j], m2[6 + j]};
    memcpy(pa, p1, 24); memcpy(pb, p2, 24);
    for (int k = 0; k < 3; k++) {
      if (k != i) { double ax[3] = {m1[k], m1[3 + k], m1[6 + k]}; double sg = dot3(bn, ax) > 0 ? 1 : -1; for (int c = 0; c < 3; c++) pa[c] += sg * s1[k] * ax[c]; }
      if (k != j) { double ax[3] = {m2[k], m2[3 + k], m2[6 + k]}; double sg = dot3(bn, ax) > 0 ? -1 : 1; for (int c = 0; c < 3; c++) pb[c] += sg * s2[k] * ax[c]; }
    }
    double w[3] = {pb[0] - pa[0], pb[1] - pa[1], pb[2] - pa[2]};
    double uaub = dot3(ua, ub), q1 = dot3(ua, w), q2 = -dot3(ub, w), den = 1 - uaub * uaub;
    double al = den > 1e-12 ? (q1 + uaub * q2) / den : 0, be = den > 1e-12 ? (uaub * q1 + q2) / den : 0;
    for (int k = 0; k < 3; k++) { pa[k] += al * ua[k]; pb[k] += be * ub[k]; }
    out[0].dist = best; memcpy(out[0].normal, bn, 24);
    for (int k = 0; k < 3; k++) out[0].pos[k] = 0.5 * (pa[k] + pb[k]);
    return 1;
  }
  /* face contact: reference box A owns the face, incident box B is clipped against its side planes */
  const double *pa = p1, *ma = m1, *sa = s1, *pb = p2, *mb = m2, *sb = s2; int ai = code, flip = 0;
  if (code >= 3) { pa = p2; ma = m2; sa = s2; pb = p1; mb = m1; sb = s1; ai = code - 3; flip = 1; }
  double nref[3] = {ma[ai], ma[3 + ai], ma[6 + ai]}, d[3] = {pb[0] - pa[0], pb[1] - pa[1], pb[2] - pa[2]};
  if (dot3(nref, d) < 0) for (int k = 0; k < 3; k++) nref[k] = -nref[k];
  int bj = 0; double bd = -1;
  for (int j = 0; j < 3; j++) { double ax[3] = {mb[j], mb[3 + j], mb[6 + j]}; double v = fabs(dot3(nref, ax)); if (v > bd) { bd = v; bj = j; } }
  double nb[3] = {mb[bj], mb[3 + bj], mb[6 + bj]};
  if (dot3(nb, nref) > 0) for (int k = 0; k < 3; k++) nb[k] = -nb[k];
  int k1 = (bj + 1) % 3, k2 = (bj + 2) % 3;
  double poly[16][3], tmp[16][3]; int np = 4;
  for (int c = 0; c < 4; c++) {
    double sg1 = (c == 0 || c == 3) ? 1 : -1, sg2 = (c < 2) ? 1 : -1;
    for (int k = 0; k < 3; k++) poly[c][k] = pb[k] + nb[k] * sb[bj] + sg1 * sb[k1] * mb[3 * k + k1] + sg2 * sb[k2] * mb[3 * k + k2];
  }
  int u1 = (ai + 1) % 3, u2 = (ai + 2) % 3;
  for (int pl = 0; pl < 4; pl++) { /* Sutherland-Hodgman against +-u1, +-u2 slabs of A */
    int u = pl < 2 ? u1 : u2; double sg = (pl & 1) ? -1 : 1;
    double ax[3] = {sg * ma[u], sg * ma[3 + u], sg * ma[6 + u]}, lim = sa[u];
    int nn = 0;
    for (int v = 0; v < np; v++) {
      const double *A = poly[v], *Bp = poly[(v + 1) % np];
      double da = (A[0] - pa[0]) * ax[0] + (A[1] - pa[1]) * ax[1] + (A[2] - pa[2]) * ax[2] - lim;
      double db = (Bp[0] - pa[0]) * ax[0] + (Bp[1] - pa[1]) * ax[1] + (Bp[2] - pa[2]) * ax[2] - lim;
      if (fabs(da) < BB_TIE) da = 0;   /* a vertex within a micron of the clip plane lies ON it (aligned boxes: gripper pads, cube on cube) */
      if (fabs(db) < BB_TIE) db = 0;
      if (da <= 0) { memcpy(tmp[nn++], A, 24); }
      if ((da < 0 && db > 0) || (da > 0 && db < 0)) { double s = da / (da - db); for (int k = 0; k < 3; k++) tmp[nn][k] = A[k] + s * (Bp[k] - A[k]); nn++; }
    }
    np = nn; memcpy(poly, tmp, sizeof(double) * 3 * np);
    if (np == 0) break;
  }
  int cnt = 0;
  for (int v = 0; v < np && cnt < 8; v++) {
    double dist = (poly[v][0] - pa[0]) * nref[0] + (poly[v][1] - pa[1]) * nref[1] + (poly[v][2] - pa[2]) * nref[2] - sa[ai];
    if (dist > margin) continue;
    out[cnt].dist = dist;
    for (int k = 0; k < 3; k++) { out[cnt].pos[k] = poly[v][k] - 0.5 * dist * nref[k]; out[cnt].normal[k] = flip ? -nref[k] : nref[k]; }
    cnt++;
  }
  return cnt;
}


/* ---- analytic pairs of the sphere / cylinder prop classes (mjc_PlaneSphere, mjc_PlaneCylinder, mjc_SphereSphere, mjc_SphereBox,
 * mjc_SphereCylinder restated from their geometry; RECOLLECTION of MuJoCo 3.2.7, unpinned like the rest).  The normal always points
 * from the pair's first geom to its second, in the PAIR TABLE's order (MuJoCo re-sorts a pair by geom type; the contact is the same
 * with normal and geom1 / geom2 swapped). */
static int plane_sphere(const double* ppos, const double* pmat, const double* c, double r, double margin, rawcon_t* out) {
  const double n[3] = {pmat[2], pmat[5], pmat[8]}, dif[3] = {c[0] - ppos[0], c[1] - ppos[1], c[2] - ppos[2]};
  const double dist = dot3(dif, n) - r;
  if (dist > margin) return 0;
  out[0].dist = dist; memcpy(out[0].normal, n, 24);
  for (int k = 0; k < 3; k++) out[0].pos[k] = c[k] - n[k] * (r + 0.5 * dist);
  return 1;
}
/* rim points of the two end discs that lie lowest along -n (a cylinder on its side rests on a line: two points), plus two more
 * points of the nearer disc 120 degrees around (a standing cylinder rests on a tripod); each kept when inside the margin */
static int plane_cylinder(const double* ppos, const double* pmat, const double* c, const double* mat, double r, double h, double margin, rawcon_t* out) {
  const double n[3] = {pmat[2], pmat[5], pmat[8]}, a[3] = {mat[2], mat[5], mat[8]};
  const double na = dot3(n, a), sg = na > 0 ? -1.0 : 1.0;
  double v[3] = {-n[0] + na * a[0], -n[1] + na * a[1], -n[2] + na * a[2]};
  if (normalize3(v) < 1e-9) { v[0] = mat[0]; v[1] = mat[3]; v[2] = mat[6]; }   /* axis along the normal: any radial direction (local x) */
  double u[3]; cross3(u, a, v);
  int cnt = 0;
  for (int q = 0; q < 4; q++) {
    const double cap = q == 1 ? -sg : sg, ca = q < 2 ? 1.0 : -0.5, sa = q < 2 ? 0.0 : (q == 2 ? 0.8660254037844387 : -0.8660254037844387);
    double pt[3];
    for (int k = 0; k < 3; k++) pt[k] = c[k] + cap * h * a[k] + r * (ca * v[k] + sa * u[k]);
    const double dif[3] = {pt[0] - ppos[0], pt[1] - ppos[1], pt[2] - ppos[2]}, dist = dot3(dif, n);
    if (dist > margin) continue;
    out[cnt].dist = dist; memcpy(out[cnt].normal, n, 24);
    for (int k = 0; k < 3; k++) out[cnt].pos[k] = pt[k] - 0.5 * dist * n[k];
    cnt++;
  }
  return cnt;
}
static int sphere_sphere(const double* c1, double r1, const double* c2, double r2, double margin, rawcon_t* out) {
  double d[3] = {c2[0] - c1[0], c2[1] - c1[1], c2[2] - c1[2]};
  const double L = normalize3(d), dist = L - r1 - r2;
  if (dist > margin) return 0;
  out[0].dist = dist; memcpy(out[0].normal, d, 24);
  for (int k = 0; k < 3; k++) out[0].pos[k] = c1[k] + d[k] * (r1 + 0.5 * dist);
  return 1;
}
/* closest point of a solid (box: half extents hs; cylinder: radius hs[0], half height hs[1], axis z) to the sphere centre, in the
 * solid's frame; a centre inside the solid is pushed out through the nearest face.  normal: from the SOLID to the sphere. */
static int sphere_solid(int solid_type, const double* cs, double rs, const double* cb, const double* mb, const double* hs, double margin, rawcon_t* out) {
  double dif[3] = {cs[0] - cb[0], cs[1] - cb[1], cs[2] - cb[2]}, lc[3], q[3], nl[3], depth;
  mulmatTvec3(lc, mb, dif);
  int inside;
  if (solid_type == B2P_GEOM_BOX) {
    inside = 1;
    for (int k = 0; k < 3; k++) { q[k] = lc[k] < -hs[k] ? -hs[k] : (lc[k] > hs[k] ? hs[k] : lc[k]); if (q[k] != lc[k]) inside = 0; }
    if (inside) {
      int best = 0; double bd = 1e300;
      for (int k = 0; k < 3; k++) { const double d = hs[k] - fabs(lc[k]); if (d < bd) { bd = d; best = k; } }
      nl[0] = nl[1] = nl[2] = 0; nl[best] = lc[best] >= 0 ? 1 : -1; q[best] = nl[best] * hs[best]; depth = -bd;
    }
  } else {
    const double rho = sqrt(lc[0] * lc[0] + lc[1] * lc[1]), r = hs[0], h = hs[1];
    inside = rho < r && fabs(lc[2]) < h;
    if (inside) {
      if (r - rho < h - fabs(lc[2])) { const double ir = rho > MINVAL ? 1 / rho : 0; nl[0] = rho > MINVAL ? lc[0] * ir : 1; nl[1] = lc[1] * ir; nl[2] = 0; q[0] = nl[0] * r; q[1] = nl[1] * r; q[2] = lc[2]; depth = -(r - rho); }
      else { nl[0] = nl[1] = 0; nl[2] = lc[2] >= 0 ? 1 : -1; q[0] = lc[0]; q[1] = lc[1]; q[2] = nl[2] * h; depth = -(h - fabs(lc[2])); }
    } else {
      const double sc = rho > r ? r / rho : 1.0;
      q[0] = lc[0] * sc; q[1] = lc[1] * sc; q[2] = lc[2] < -h ? -h : (lc[2] > h ? h : lc[2]);
    }
  }
  if (!inside) { for (int k = 0; k < 3; k++) nl[k] = lc[k] - q[k]; depth = normalize3(nl); }
  const double dist = depth - rs;
  if (dist > margin) return 0;
  double nw[3], qw[3];
  mulmatvec3(nw, mb, nl); mulmatvec3(qw, mb, q);
  out[0].dist = dist; memcpy(out[0].normal, nw, 24);
  for (int k = 0; k < 3; k++) out[0].pos[k] = cb[k] + qw[k] + 0.5 * dist * nw[k];
  return 1;
}
static void mix_params(orc_t* o, int g1, int g2, contact_t* c) {
  /* App. B item 10: priority wins; equal priority -> max friction, solmix-weighted solref/solimp, max condim */
  const int *prio = FI(o, B2PM_F_GEOM_PRIORITY), *condim = FI(o, B2PM_F_GEOM_CONDIM);
  const double *fr = FD(o, B2PM_F_GEOM_FRICTION), *sr = FD(o, B2PM_F_GEOM_SOLREF), *si = FD(o, B2PM_F_GEOM_SOLIMP), *sm = FD(o, B2PM_F_GEOM_SOLMIX);
  double f[3];
  if (prio[g1] != prio[g2]) {
    int g = prio[g1] > prio[g2] ? g1 : g2;
    c->dim = condim[g]; memcpy(f, fr + 3 * g, 24); memcpy(c->solref, sr + 2 * g, 16); memcpy(c->solimp, si + 5 * g, 40);
  } else {
    c->dim = condim[g1] > condim[g2] ? condim[g1] : condim[g2];
    for (int k = 0; k < 3; k++) f[k] = fr[3 * g1 + k] > fr[3 * g2 + k] ? fr[3 * g1 + k] : fr[3 * g2 + k];
    double mix;
    if (sm[g1] >= MINVAL && sm[g2] >= MINVAL) mix = sm[g1] / (sm[g1] + sm[g2]);
    else if (sm[g1] < MINVAL && sm[g2] < MINVAL) mix = 0.5; else mix = sm[g1] < MINVAL ? 0 : 1;
    if (sr[2 * g1] > 0 && sr[2 * g2] > 0) for (int k = 0; k < 2; k++) c->solref[k] = mix * sr[2 * g1 + k] + (1 - mix) * sr[2 * g2 + k];
    else for (int k = 0; k < 2; k++) c->solref[k] = sr[2 * g1 + k] < sr[2 * g2 + k] ? sr[2 * g1 + k] : sr[2 * g2 + k];
    for (int k = 0; k < 5; k++) c->solimp[k] = mix * si[5 * g1 + k] + (1 - mix) * si[5 * g2 + k];
  }
  c->friction[0] = c->friction[1] = f[0]; c->friction[2] = f[1]; c->friction[3] = c->friction[4] = f[2];
}

static void collision(orc_t* o) {
  const int *pg1 = FI(o, B2PM_F_PAIR_GEOM1), *pg2 = FI(o, B2PM_F_PAIR_GEOM2), *gtype = FI(o, B2PM_F_GEOM_TYPE);
  const int *gprop = FI(o, B2PM_F_GEOM_PROPID), *gmesh = FI(o, B2PM_F_GEOM_MESHID), *vadr = FI(o, B2PM_F_MESH_VERTADR), *vnum = FI(o, B2PM_F_MESH_VERTNUM);
  const double *gmargin = FD(o, B2PM_F_GEOM_MARGIN), *ggap = FD(o, B2PM_F_GEOM_GAP), *mvert = FD(o, B2PM_F_MESH_VERT);
  o->ncon = 0;
  for (int p = 0; p < o->npair; p++) {
    int g1 = pg1[p], g2 = pg2[p];
    if ((gprop[g1] >= 0 && !o->prop_active[gprop[g1]]) || (gprop[g2] >= 0 && !o->prop_active[gprop[g2]])) continue;
    double margin = gmargin[g1] > gmargin[g2] ? gmargin[g1] : gmargin[g2], gap = ggap[g1] > ggap[g2] ? ggap[g1] : ggap[g2];
    double md = o->full_contacts ? margin : margin - gap;
    const double *x1 = o->geom_xpos + 3 * g1, *x2 = o->geom_xpos + 3 * g2, *m1 = o->geom_xmat + 9 * g1, *m2 = o->geom_xmat + 9 * g2;
    int t1 = gtype[g1], t2 = gtype[g2];
    rawcon_t raw[8]; int n = 0;
    if (t1 == B2P_GEOM_PLANE) {
      double nrm[3] = {m1[2], m1[5], m1[8]}, dif[3] = {x2[0] - x1[0], x2[1] - x1[1], x2[2] - x1[2]};
      if (dot3(dif, nrm) - o->geom_rbound[g2] > md) continue;
      if (t2 == B2P_GEOM_BOX) n = plane_box(x1, m1, x2, m2, o->geom_size + 3 * g2, md, raw);
      else if (t2 == B2P_GEOM_SPHERE) n = plane_sphere(x1, m1, x2, o->geom_size[3 * g2], md, raw);
      else if (t2 == B2P_GEOM_CYLINDER) n = plane_cylinder(x1, m1, x2, m2, o->geom_size[3 * g2], o->geom_size[3 * g2 + 1], md, raw);
      else if (t2 == B2P_GEOM_MESH) {   /* an arm link below the table edge reaches the ground */
        int v0 = vadr[gmesh[g2]];
        cobj_t h = {x2, m2, t2, o->geom_size + 3 * g2, mvert + 3 * v0, vnum[gmesh[g2]], 0, FI(o, B2PM_F_MESH_NBRADR) + v0, FI(o, B2PM_F_MESH_NBRNUM) + v0, FI(o, B2PM_F_MESH_NBR), 0};
        n = plane_convex(x1, m1, &h, md, raw);
      } else continue;
    } else {
      double dif[3] = {x2[0] - x1[0], x2[1] - x1[1], x2[2] - x1[2]};
      if (norm3(dif) > o->geom_rbound[g1] + o->geom_rbound[g2] + md) continue;
#ifdef ORC_FAST
      { /* mid phase of the timing build: the geoms' oriented bounding boxes along their 6 face normals (conservative: never drops a contact) */
        const double* ab = FD(o, B2PM_F_GEOM_AABB);
        const double *c1 = ab + 6 * g1, *c2 = ab + 6 * g2, *h1 = o->geom_half + 3 * g1, *h2 = o->geom_half + 3 * g2;
        double w1[3], w2[3], t[3], sep = 0; mulmatvec3(w1, m1, c1); mulmatvec3(w2, m2, c2);
        for (int k = 0; k < 3; k++) t[k] = (x2[k] + w2[k]) - (x1[k] + w1[k]);
        for (int i = 0; i < 3 && !sep; i++) {
          double ti = t[0] * m1[i] + t[1] * m1[3 + i] + t[2] * m1[6 + i], r = h1[i];
          for (int j = 0; j < 3; j++) r += h2[j] * fabs(m1[i] * m2[j] + m1[3 + i] * m2[3 + j] + m1[6 + i] * m2[6 + j]);
          if (fabs(ti) > r + md + 1e-6) sep = 1;
        }
        for (int j = 0; j < 3 && !sep; j++) {
          double tj = t[0] * m2[j] + t[1] * m2[3 + j] + t[2] * m2[6 + j], r = h2[j];
          for (int i = 0; i < 3; i++) r += h1[i] * fabs(m1[i] * m2[j] + m1[3 + i] * m2[3 + j] + m1[6 + i] * m2[6 + j]);
          if (fabs(tj) > r + md + 1e-6) sep = 1;
        }
        if (sep) continue;
      }
#endif
      int convex = 1;
      if (t1 == B2P_GEOM_BOX && t2 == B2P_GEOM_BOX) {
        n = box_box(x1, m1, o->geom_size + 3 * g1, x2, m2, o->geom_size + 3 * g2, md, raw);
        convex = (n < 0); if (n < 0) n = 0;
      } else if (t1 == B2P_GEOM_SPHERE && t2 == B2P_GEOM_SPHERE) { n = sphere_sphere(x1, o->geom_size[3 * g1], x2, o->geom_size[3 * g2], md, raw); convex = 0; }
      else if (t1 == B2P_GEOM_SPHERE && (t2 == B2P_GEOM_BOX || t2 == B2P_GEOM_CYLINDER)) {     /* normal solid -> sphere = second -> first: flip */
        n = sphere_solid(t2, x1, o->geom_size[3 * g1], x2, m2, o->geom_size + 3 * g2, md, raw); convex = 0;
        for (int i = 0; i < n; i++) for (int k = 0; k < 3; k++) raw[i].normal[k] = -raw[i].normal[k];
      } else if (t2 == B2P_GEOM_SPHERE && (t1 == B2P_GEOM_BOX || t1 == B2P_GEOM_CYLINDER)) { n = sphere_solid(t1, x2, o->geom_size[3 * g2], x1, m1, o->geom_size + 3 * g1, md, raw); convex = 0; }
      if (convex) {
        cobj_t a = {x1, m1, t1, o->geom_size + 3 * g1, NULL, 0, md, NULL, NULL, NULL, 0}, b = {x2, m2, t2, o->geom_size + 3 * g2, NULL, 0, md, NULL, NULL, NULL, 0};
        const int *nadr = FI(o, B2PM_F_MESH_NBRADR), *nnum = FI(o, B2PM_F_MESH_NBRNUM), *nbr = FI(o, B2PM_F_MESH_NBR);
        if (t1 == B2P_GEOM_MESH) { int v0 = vadr[gmesh[g1]]; a.vert = mvert + 3 * v0; a.nvert = vnum[gmesh[g1]]; a.nbradr = nadr + v0; a.nbrnum = nnum + v0; a.nbr = nbr; }
        if (t2 == B2P_GEOM_MESH) { int v0 = vadr[gmesh[g2]]; b.vert = mvert + 3 * v0; b.nvert = vnum[gmesh[g2]]; b.nbradr = nadr + v0; b.nbrnum = nnum + v0; b.nbr = nbr; }
        double depth, dir[3], pos[3];
        if (mpr_penetration(&a, &b, o->ccd_iterations, o->ccd_tolerance, &depth, dir, pos) && depth >= 0) {
          raw[0].dist = md - depth; memcpy(raw[0].normal, dir, 24); memcpy(raw[0].pos, pos, 24); n = 1;
        }
      }
    }
    for (int i = 0; i < n && o->ncon < ORC_MAXCON; i++) {
      if (!(raw[i].dist < md)) continue;
      contact_t* c = &o->con[o->ncon++];
      memset(c, 0, sizeof *c);
      c->dist = raw[i].dist; memcpy(c->pos, raw[i].pos, 24); memcpy(c->frame, raw[i].normal, 24);
      make_frame(c->frame);
      c->includemargin = margin - gap; c->geom1 = g1; c->geom2 = g2;
      mix_params(o, g1, g2, c);
      c->exclude = !(c->dist < c->includemargin);
      c->efc_address = -1;
    }
  }
}

/* ------------------------------------------------------------------ constraints (App. B items 8, 9) */
static double impedance(const double* si, double pos, double margin) {
  double d0 = fmin(MAXIMP, fmax(MINIMP, si[0])), d1 = fmin(MAXIMP, fmax(MINIMP, si[1])), w = fmax(0, si[2]);
  double mid = fmin(MAXIMP, fmax(MINIMP, si[3])), pw = fmax(1, si[4]);
  if (d0 == d1 || w <= MINVAL) return 0.5 * (d0 + d1);
  double x = fabs((pos - margin) / w), y;
  if (x >= 1) return d1;
  if (x <= 0) return d0;
  if (pw == 1) y = x;
  else if (x <= mid) y = pow(x, pw) / pow(mid, pw - 1);
  else y = 1 - pow(1 - x, pw) / pow(1 - mid, pw - 1);
  return d0 + y * (d1 - d0);
}
static int add_row(orc_t* o, int type, int id, double pos, double margin, double diag) {
  int r = o->nefc;
  if (r >= ORC_MAXEFC) return -1;
  o->nefc++;
  memset(o->efc_J + (size_t)r * o->nv, 0, 8 * o->nv);
  o->efc_type[r] = type; o->efc_id[r] = id; o->efc_pos[r] = pos; o->efc_margin[r] = margin; o->efc_diagApprox[r] = diag;
  return r;
}
static void finish_row(orc_t* o, int r, const double* solref_in, const double* solimp, int friction_row) {
  double solref[2] = {solref_in[0], solref_in[1]};
  if (solref[0] > 0 && solref[0] < 2 * o->timestep) solref[0] = 2 * o->timestep; /* refsafe */
  double imp = impedance(solimp, o->efc_pos[r], o->efc_margin[r]);
  double dmax = fmin(MAXIMP, fmax(MINIMP, solimp[1])), K, Bd;
  if (solref[0] > 0) { K = 1 / fmax(MINVAL, dmax * dmax * solref[0] * solref[0] * solref[1] * solref[1]); Bd = 2 / fmax(MINVAL, dmax * solref[0]); }
  else { K = -solref[0] / fmax(MINVAL, dmax * dmax); Bd = -solref[1] / fmax(MINVAL, dmax); }
  if (friction_row) K = 0;
  o->efc_KBIP[4 * r] = K; o->efc_KBIP[4 * r + 1] = Bd; o->efc_KBIP[4 * r + 2] = imp;
  o->efc_R[r] = fmax(MINVAL, (1 - imp) / imp * o->efc_diagApprox[r]);
}
static void make_constraint(orc_t* o) {
  int nv = o->nv;
  double* jp1 = o->jacbuf; double *jp2 = jp1 + 3 * nv, *jr1 = jp2 + 3 * nv, *jr2 = jr1 + 3 * nv;
  o->nefc = 0;
  /* equality */
  const int *et = FI(o, B2PM_F_EQ_TYPE), *e1 = FI(o, B2PM_F_EQ_OBJ1ID), *e2 = FI(o, B2PM_F_EQ_OBJ2ID);
  const double *ed = FD(o, B2PM_F_EQ_DATA), *esr = FD(o, B2PM_F_EQ_SOLREF), *esi = FD(o, B2PM_F_EQ_SOLIMP), *qpos0 = FD(o, B2PM_F_QPOS0);
  const int *jq = FI(o, B2PM_F_JNT_QPOSADR), *jd = FI(o, B2PM_F_JNT_DOFADR);
  for (int e = 0; e < o->neq; e++) {
    if (et[e] == B2P_EQ_CONNECT) {
      int b1 = e1[e], b2 = e2[e]; double g1[3], g2[3];
      mulmatvec3(g1, o->xmat + 9 * b1, ed + 11 * e); mulmatvec3(g2, o->xmat + 9 * b2, ed + 11 * e + 3);
      for (int k = 0; k < 3; k++) { g1[k] += o->xpos[3 * b1 + k]; g2[k] += o->xpos[3 * b2 + k]; }
      jac_point(o, b1, g1, jp1, NULL); jac_point(o, b2, g2, jp2, NULL);
      double diag = o->body_invweight0[2 * b1] + o->body_invweight0[2 * b2];
      for (int k = 0; k < 3; k++) {
        int r = add_row(o, ROW_EQ, e, g1[k] - g2[k], 0, diag);
        for (int i = 0; i < nv; i++) o->efc_J[(size_t)r * nv + i] = jp1[k * nv + i] - jp2[k * nv + i];
      }
      /* MuJoCo computes the impedance of a connect from the norm of the 3-vector residual */
      double nrm = sqrt((g1[0] - g2[0]) * (g1[0] - g2[0]) + (g1[1] - g2[1]) * (g1[1] - g2[1]) + (g1[2] - g2[2]) * (g1[2] - g2[2]));
      for (int k = 0; k < 3; k++) { int r = o->nefc - 3 + k; double keep = o->efc_pos[r]; o->efc_pos[r] = nrm; finish_row(o, r, esr + 2 * e, esi + 5 * e, 0); o->efc_pos[r] = keep; }
    } else {
      int j1 = e1[e], j2 = e2[e]; const double* c = ed + 11 * e;
      double x = o->qpos[jq[j2]] - qpos0[jq[j2]], pos = o->qpos[jq[j1]] - qpos0[jq[j1]];
      double poly = c[0] + x * (c[1] + x * (c[2] + x * (c[3] + x * c[4]))), der = c[1] + x * (2 * c[2] + x * (3 * c[3] + x * 4 * c[4]));
      int r = add_row(o, ROW_EQ, e, pos - poly, 0, o->dof_invweight0[jd[j1]] + o->dof_invweight0[jd[j2]]);
      o->efc_J[(size_t)r * nv + jd[j1]] = 1; o->efc_J[(size_t)r * nv + jd[j2]] = -der;
      finish_row(o, r, esr + 2 * e, esi + 5 * e, 0);
    }
  }
  /* joint limits */
  const int *jt = FI(o, B2PM_F_JNT_TYPE), *jl = FI(o, B2PM_F_JNT_LIMITED);
  const double *jr = FD(o, B2PM_F_JNT_RANGE), *jm = FD(o, B2PM_F_JNT_MARGIN), *jsr = FD(o, B2PM_F_JNT_SOLREF), *jsi = FD(o, B2PM_F_JNT_SOLIMP);
  for (int j = 0; j < o->njnt; j++) {
    if (jt[j] != B2P_JNT_HINGE || !jl[j]) continue;
    double v = o->qpos[jq[j]];
    for (int side = -1; side <= 1; side += 2) {
      double dist = side * (jr[2 * j + (side + 1) / 2] - v);
      if (dist < jm[j]) {
        int r = add_row(o, ROW_LIMIT, j, dist, jm[j], o->dof_invweight0[jd[j]]);
        o->efc_J[(size_t)r * nv + jd[j]] = -side;
        finish_row(o, r, jsr + 2 * j, jsi + 5 * j, 0);
      }
    }
  }
  /* contacts (elliptic cones; pyramidal below) */
  const int* gb = FI(o, B2PM_F_GEOM_BODYID);
  for (int ci = 0; ci < o->ncon; ci++) {
    contact_t* c = &o->con[ci];
    if (c->exclude) continue;
    int b1 = gb[c->geom1], b2 = gb[c->geom2];
    jac_point(o, b1, c->pos, jp1, jr1); jac_point(o, b2, c->pos, jp2, jr2);
    double tran = o->geom_invweight0[2 * c->geom1] + o->geom_invweight0[2 * c->geom2];
    double rot = o->geom_invweight0[2 * c->geom1 + 1] + o->geom_invweight0[2 * c->geom2 + 1];
    c->efc_address = o->nefc;
    if (o->cone == B2P_CONE_PYRAMIDAL && c->dim > 1) {
      /* MuJoCo's pyramidal cone (mj_instantiateContact / mj_makeImpedance, restated from the published algorithm): 2 (dim - 1) scalar
       * inequality rows per contact, one per pyramid edge, J = J_normal +- mu_k J_tangent_k; every edge carries the contact's
       * distance (so K, B and the impedance are the normal row's); diagApprox = tran + mu_k^2 (tran | rot); and after the impedance
       * pass all edges share R_py = 2 mu_reg^2 R_first with mu_reg = mu_1 / sqrt(impratio).  In the solver an edge is an ordinary
       * inequality row (force >= 0), which is what ROW_LIMIT means there. */
      if (o->nefc + 2 * (c->dim - 1) > ORC_MAXEFC) { c->efc_address = -1; c->exclude = 2; continue; }
      int r0 = o->nefc;
      for (int k = 1; k < c->dim; k++) for (int sgn = 1; sgn >= -1; sgn -= 2) {
        const double mu = c->friction[k - 1];
        int r = add_row(o, ROW_LIMIT, ci, c->dist, c->includemargin, tran + mu * mu * (k < 3 ? tran : rot));
        const double* n = c->frame; const double* ax = c->frame + 3 * (k % 3);
        const double *ja = k < 3 ? jp1 : jr1, *jb = k < 3 ? jp2 : jr2;
        for (int i = 0; i < nv; i++)
          o->efc_J[(size_t)r * nv + i] = n[0] * (jp2[i] - jp1[i]) + n[1] * (jp2[nv + i] - jp1[nv + i]) + n[2] * (jp2[2 * nv + i] - jp1[2 * nv + i])
                                         + sgn * mu * (ax[0] * (jb[i] - ja[i]) + ax[1] * (jb[nv + i] - ja[nv + i]) + ax[2] * (jb[2 * nv + i] - ja[2 * nv + i]));
        finish_row(o, r, c->solref, c->solimp, 0);
      }
      { const double R1 = o->efc_R[r0] / fmax(MINVAL, o->impratio); c->mu = c->friction[0] * sqrt(R1 / o->efc_R[r0]);
        const double Rpy = 2 * c->mu * c->mu * o->efc_R[r0];
        for (int r = r0; r < o->nefc; r++) o->efc_R[r] = Rpy; }
      continue;
    }
    int r0 = -1;
    for (int k = 0; k < c->dim; k++) {
      int r = add_row(o, ROW_CONTACT, ci, k == 0 ? c->dist : 0, c->includemargin, k < 3 ? tran : rot);
      if (r < 0) { c->efc_address = -1; c->exclude = 2; break; }
      if (k == 0) r0 = r;
      const double* ax = c->frame + 3 * (k % 3);
      const double *ja = k < 3 ? jp1 : jr1, *jb = k < 3 ? jp2 : jr2;
      for (int i = 0; i < nv; i++) o->efc_J[(size_t)r * nv + i] = ax[0] * (jb[i] - ja[i]) + ax[1] * (jb[nv + i] - ja[nv + i]) + ax[2] * (jb[2 * nv + i] - ja[2 * nv + i]);
      finish_row(o, r, c->solref, c->solimp, k > 0);
    }
    if (r0 >= 0 && c->dim > 1 && c->efc_address >= 0) {
      /* elliptic friction regularisation (App. B item 9) */
      o->efc_R[r0 + 1] = o->efc_R[r0] / fmax(MINVAL, o->impratio);
      for (int k = 2; k < c->dim; k++) o->efc_R[r0 + k] = o->efc_R[r0 + 1] * c->friction[0] * c->friction[0] / (c->friction[k - 1] * c->friction[k - 1]);
      c->mu = c->friction[0] * sqrt(o->efc_R[r0 + 1] / o->efc_R[r0]);
    }
  }
  for (int r = 0; r < o->nefc; r++) {
    o->efc_D[r] = 1 / o->efc_R[r];
    const double* J = o->efc_J + (size_t)r * nv; unsigned char* ind = o->efc_ind + (size_t)r * MAXNV; int n = 0;
    for (int i = 0; i < nv; i++) if (J[i] != 0) ind[n++] = (unsigned char)i;
    o->efc_nnz[r] = n;
  }
}
/* J_r . x over the row's non-zero entries (same summation order as the dense loop: skipped terms are exact zeros) */
static double row_dot(const orc_t* o, int r, const double* x) {
  const double* J = o->efc_J + (size_t)r * o->nv; const unsigned char* ind = o->efc_ind + (size_t)r * MAXNV; double v = 0;
  for (int k = 0; k < o->efc_nnz[r]; k++) v += J[ind[k]] * x[ind[k]];
  return v;
}
/* mj_referenceConstraint: aref = -B*vel - K*imp*(pos-margin) */
static void reference_constraint(orc_t* o) {
  int nv = o->nv;
  for (int r = 0; r < o->nefc; r++) {
    double v = row_dot(o, r, o->qvel);
    o->efc_vel[r] = v;
    o->efc_aref[r] = -o->efc_KBIP[4 * r + 1] * v - o->efc_KBIP[4 * r] * o->efc_KBIP[4 * r + 2] * (o->efc_pos[r] - o->efc_margin[r]);
  }
}

/* ------------------------------------------------------------------ velocity stage: mj_comVel, mj_passive, mj_rne */
static void com_vel(orc_t* o) {
  const int *par = FI(o, B2PM_F_BODY_PARENTID), *jnum = FI(o, B2PM_F_BODY_JNTNUM), *jadr = FI(o, B2PM_F_BODY_JNTADR);
  const int *jtype = FI(o, B2PM_F_JNT_TYPE), *jdof = FI(o, B2PM_F_JNT_DOFADR);
  memset(o->cvel, 0, 48);
  for (int i = 1; i < o->nbody; i++) {
    double v[6]; memcpy(v, o->cvel + 6 * par[i], 48);
    for (int jj = 0; jj < jnum[i]; jj++) {
      int j = jadr[i] + jj, d = jdof[j];
      if (jtype[j] == B2P_JNT_FREE) {
        memset(o->cdof_dot + 6 * d, 0, 18 * 8);
        for (int k = 0; k < 3; k++) for (int c = 0; c < 6; c++) v[c] += o->cdof[6 * (d + k) + c] * o->qvel[d + k];
        for (int k = 3; k < 6; k++) cross_motion(o->cdof_dot + 6 * (d + k), v, o->cdof + 6 * (d + k));
        for (int k = 3; k < 6; k++) for (int c = 0; c < 6; c++) v[c] += o->cdof[6 * (d + k) + c] * o->qvel[d + k];
      } else {
        cross_motion(o->cdof_dot + 6 * d, v, o->cdof + 6 * d);
        for (int c = 0; c < 6; c++) v[c] += o->cdof[6 * d + c] * o->qvel[d];
      }
    }
    memcpy(o->cvel + 6 * i, v, 48);
  }
}
static void passive(orc_t* o) {
  const double *damp = FD(o, B2PM_F_DOF_DAMPING), *stiff = FD(o, B2PM_F_JNT_STIFFNESS), *qs = FD(o, B2PM_F_QPOS_SPRING);
  const int *jt = FI(o, B2PM_F_JNT_TYPE), *jq = FI(o, B2PM_F_JNT_QPOSADR), *jd = FI(o, B2PM_F_JNT_DOFADR);
  for (int i = 0; i < o->nv; i++) o->qfrc_passive[i] = -damp[i] * o->qvel[i];
  for (int j = 0; j < o->njnt; j++) if (jt[j] == B2P_JNT_HINGE && stiff[j] != 0) o->qfrc_passive[jd[j]] -= stiff[j] * (o->qpos[jq[j]] - qs[jq[j]]);
}
static void rne_bias(orc_t* o) {
  const int *par = FI(o, B2PM_F_BODY_PARENTID), *dnum = FI(o, B2PM_F_BODY_DOFNUM), *dadr = FI(o, B2PM_F_BODY_DOFADR), *dbody = FI(o, B2PM_F_DOF_BODYID);
  double cacc[6 * MAXBODY], cfrc[6 * MAXBODY];
  memset(cacc, 0, 48); cacc[3] = -o->gravity[0]; cacc[4] = -o->gravity[1]; cacc[5] = -o->gravity[2];
  memset(cfrc, 0, 48);
  for (int i = 1; i < o->nbody; i++) {
    double* a = cacc + 6 * i; memcpy(a, cacc + 6 * par[i], 48);
    for (int k = 0; k < dnum[i]; k++) { int d = dadr[i] + k; for (int c = 0; c < 6; c++) a[c] += o->cdof_dot[6 * d + c] * o->qvel[d]; }
    double Ia[6], Iv[6], cf[6];
    mul_inert_vec(Ia, o->cinert + 10 * i, a); mul_inert_vec(Iv, o->cinert + 10 * i, o->cvel + 6 * i);
    cross_force(cf, o->cvel + 6 * i, Iv);
    for (int c = 0; c < 6; c++) cfrc[6 * i + c] = Ia[c] + cf[c];
  }
  for (int i = o->nbody - 1; i > 0; i--) if (par[i] > 0) for (int c = 0; c < 6; c++) cfrc[6 * par[i] + c] += cfrc[6 * i + c];
  for (int d = 0; d < o->nv; d++) { double v = 0; for (int c = 0; c < 6; c++) v += o->cdof[6 * d + c] * cfrc[6 * dbody[d] + c]; o->qfrc_bias[d] = v; }
}

/* ------------------------------------------------------------------ mj_fwdActuation (App. B item 7) */
static void actuation(orc_t* o) {
  const int *tn = FI(o, B2PM_F_ACT_TRNNUM), *td = FI(o, B2PM_F_ACT_TRNDOF), *cl = FI(o, B2PM_F_ACT_CTRLLIMITED), *fl = FI(o, B2PM_F_ACT_FORCELIMITED);
  const double *tc = FD(o, B2PM_F_ACT_TRNCOEF), *gain = FD(o, B2PM_F_ACT_GAIN), *bias = FD(o, B2PM_F_ACT_BIAS), *cr = FD(o, B2PM_F_ACT_CTRLRANGE), *fr = FD(o, B2PM_F_ACT_FORCERANGE);
  const int *jq = FI(o, B2PM_F_JNT_QPOSADR), *dj = FI(o, B2PM_F_DOF_JNTID);
  memset(o->qfrc_actuator, 0, 8 * o->nv);
  for (int a = 0; a < o->nu; a++) {
    double len = 0, vel = 0;
    for (int k = 0; k < tn[a]; k++) { int d = td[2 * a + k]; len += tc[2 * a + k] * o->qpos[jq[dj[d]]]; vel += tc[2 * a + k] * o->qvel[d]; }
    o->act_length[a] = len; o->act_velocity[a] = vel;
    double c = o->ctrl[a];
    if (cl[a]) c = fmin(cr[2 * a + 1], fmax(cr[2 * a], c));
    double f = gain[a] * c + bias[3 * a] + bias[3 * a + 1] * len + bias[3 * a + 2] * vel;
    if (fl[a]) f = fmin(fr[2 * a + 1], fmax(fr[2 * a], f));
    o->act_force[a] = f;
    for (int k = 0; k < tn[a]; k++) o->qfrc_actuator[td[2 * a + k]] += tc[2 * a + k] * f;
  }
}

/* ------------------------------------------------------------------ Newton solver (App. B item 13) */
typedef struct {
  orc_t* o; double *Jaref, *Jv, *Ma, *Mv, *search, *grad, *Mgrad, *H;
  double quadGauss[3], cost, scale;
} solver_t;

static void mul_M(orc_t* o, double* r, const double* v) {
  int nv = o->nv;
  for (int i = 0; i < nv; i++) { double s = 0; for (int j = 0; j < nv; j++) s += o->qM[i * nv + j] * v[j]; r[i] = s; }
}
/* mj_constraintUpdate: forces, states, constraint cost; optional Hessian cone blocks */
static double constraint_update(orc_t* o, const double* jar, double* H) {
  double cost = 0; int nv = o->nv;
  for (int r = 0; r < o->nefc; r++) {
    if (o->efc_type[r] == ROW_EQ) { o->efc_force[r] = -o->efc_D[r] * jar[r]; o->efc_state[r] = ST_QUADRATIC; cost += 0.5 * o->efc_D[r] * jar[r] * jar[r]; }
    else if (o->efc_type[r] == ROW_LIMIT) {
      if (jar[r] < 0) { o->efc_force[r] = -o->efc_D[r] * jar[r]; o->efc_state[r] = ST_QUADRATIC; cost += 0.5 * o->efc_D[r] * jar[r] * jar[r]; }
      else { o->efc_force[r] = 0; o->efc_state[r] = ST_SATISFIED; }
    } else {
      contact_t* c = &o->con[o->efc_id[r]]; int dim = c->dim; double mu = c->mu, U[6];
      if (dim == 1) { /* frictionless */
        if (jar[r] < 0) { o->efc_force[r] = -o->efc_D[r] * jar[r]; o->efc_state[r] = ST_QUADRATIC; cost += 0.5 * o->efc_D[r] * jar[r] * jar[r]; }
        else { o->efc_force[r] = 0; o->efc_state[r] = ST_SATISFIED; }
        continue;
      }
      U[0] = jar[r] * mu;
      double T = 0;
      for (int k = 1; k < dim; k++) { U[k] = jar[r + k] * c->friction[k - 1]; T += U[k] * U[k]; }
      T = sqrt(T); double N = U[0];
      if (N >= mu * T || (T <= 0 && N >= 0)) { for (int k = 0; k < dim; k++) { o->efc_force[r + k] = 0; o->efc_state[r + k] = ST_SATISFIED; } }
      else if (mu * N + T <= 0 || (T <= 0 && N < 0)) {
        for (int k = 0; k < dim; k++) { o->efc_force[r + k] = -o->efc_D[r + k] * jar[r + k]; o->efc_state[r + k] = ST_QUADRATIC; cost += 0.5 * o->efc_D[r + k] * jar[r + k] * jar[r + k]; }
      } else {
        double Dm = o->efc_D[r] / fmax(MINVAL, mu * mu * (1 + mu * mu)), NmT = N - mu * T;
        cost += 0.5 * Dm * NmT * NmT;
        o->efc_force[r] = -Dm * NmT * mu;
        for (int k = 1; k < dim; k++) o->efc_force[r + k] = -o->efc_force[r] / T * U[k] * c->friction[k - 1];
        for (int k = 0; k < dim; k++) o->efc_state[r + k] = ST_CONE;
        if (H) { /* J_c' * (d2s/djar2) * J_c */
          double hl[36], sc[6];
          sc[0] = mu; for (int k = 1; k < dim; k++) sc[k] = c->friction[k - 1];
          hl[0] = Dm;
          for (int k = 1; k < dim; k++) hl[k] = hl[k * dim] = -Dm * mu * U[k] / T;
          for (int a = 1; a < dim; a++) for (int b = 1; b < dim; b++)
            hl[a * dim + b] = Dm * mu * mu * U[a] * U[b] / (T * T) - Dm * mu * NmT * ((a == b ? 1.0 / T : 0.0) - U[a] * U[b] / (T * T * T));
          for (int a = 0; a < dim; a++) for (int b = 0; b < dim; b++) hl[a * dim + b] *= sc[a] * sc[b];
          for (int a = 0; a < dim; a++) for (int b = 0; b < dim; b++) {
            double h = hl[a * dim + b]; if (h == 0) continue;
            const double *Ja = o->efc_J + (size_t)(r + a) * nv, *Jb = o->efc_J + (size_t)(r + b) * nv;
            const unsigned char *ia = o->efc_ind + (size_t)(r + a) * MAXNV, *ib = o->efc_ind + (size_t)(r + b) * MAXNV;
            for (int x = 0; x < o->efc_nnz[r + a]; x++) { int i = ia[x]; double t = h * Ja[i]; double* Hi = H + i * nv; for (int y = 0; y < o->efc_nnz[r + b]; y++) Hi[ib[y]] += t * Jb[ib[y]]; }
          }
        }
      }
      r += dim - 1;
    }
  }
  return cost;
}
static int cholesky(double* A, int n) { /* lower, in place */
  for (int j = 0; j < n; j++) {
    double d = A[j * n + j];
    for (int k = 0; k < j; k++) d -= A[j * n + k] * A[j * n + k];
    if (d < MINVAL) return -1;
    d = sqrt(d); A[j * n + j] = d;
    for (int i = j + 1; i < n; i++) { double s = A[i * n + j]; for (int k = 0; k < j; k++) s -= A[i * n + k] * A[j * n + k]; A[i * n + j] = s / d; }
  }
  return 0;
}
static void chol_solve(const double* L, int n, double* x) {
  for (int i = 0; i < n; i++) { double s = x[i]; for (int k = 0; k < i; k++) s -= L[i * n + k] * x[k]; x[i] = s / L[i * n + i]; }
  for (int i = n - 1; i >= 0; i--) { double s = x[i]; for (int k = i + 1; k < n; k++) s -= L[k * n + i] * x[k]; x[i] = s / L[i * n + i]; }
}
/* full cost / gradient / Newton direction at the current qacc */
static void solver_update(solver_t* s, const double* qacc) {
  orc_t* o = s->o; int nv = o->nv;
  memcpy(s->H, o->qM, 8 * nv * nv);
  double cost = constraint_update(o, s->Jaref, s->H);
  for (int r = 0; r < o->nefc; r++) if (o->efc_state[r] == ST_QUADRATIC) {
    const double* J = o->efc_J + (size_t)r * nv; double D = o->efc_D[r]; const unsigned char* ind = o->efc_ind + (size_t)r * MAXNV; int nz = o->efc_nnz[r];
    for (int a = 0; a < nz; a++) { int i = ind[a]; double t = D * J[i]; double* Hi = s->H + i * nv; for (int b = 0; b < nz; b++) Hi[ind[b]] += t * J[ind[b]]; }
  }
  memset(o->qfrc_constraint, 0, 8 * nv);
  for (int r = 0; r < o->nefc; r++) if (o->efc_force[r] != 0) { const double* J = o->efc_J + (size_t)r * nv; const unsigned char* ind = o->efc_ind + (size_t)r * MAXNV; for (int k = 0; k < o->efc_nnz[r]; k++) o->qfrc_constraint[ind[k]] += J[ind[k]] * o->efc_force[r]; }
  double gauss = 0;
  for (int i = 0; i < nv; i++) { gauss += 0.5 * (s->Ma[i] - o->qfrc_smooth[i]) * (qacc[i] - o->qacc_smooth[i]); s->grad[i] = s->Ma[i] - o->qfrc_smooth[i] - o->qfrc_constraint[i]; }
  s->cost = cost + gauss;
  if (cholesky(s->H, nv) != 0) { /* should not happen (H = M + PSD); fall back to M */
    memcpy(s->H, o->qM, 8 * nv * nv); cholesky(s->H, nv);
  }
  memcpy(s->Mgrad, s->grad, 8 * nv); chol_solve(s->H, nv, s->Mgrad);
}
typedef struct { double alpha, cost, d1, d2; } lspt_t;
static lspt_t ls_eval(solver_t* s, double a) {
  orc_t* o = s->o; lspt_t p = {a, 0, 0, 0};
  double q0 = s->quadGauss[0], q1 = s->quadGauss[1], q2 = s->quadGauss[2];
  for (int r = 0; r < o->nefc; r++) {
    double D = o->efc_D[r], j0 = s->Jaref[r], jv = s->Jv[r];
    if (o->efc_type[r] == ROW_EQ) { q0 += 0.5 * D * j0 * j0; q1 += D * j0 * jv; q2 += 0.5 * D * jv * jv; }
    else if (o->efc_type[r] == ROW_LIMIT || o->con[o->efc_id[r]].dim == 1) { if (j0 + a * jv < 0) { q0 += 0.5 * D * j0 * j0; q1 += D * j0 * jv; q2 += 0.5 * D * jv * jv; } }
    else {
      contact_t* c = &o->con[o->efc_id[r]]; int dim = c->dim; double mu = c->mu;
      double U0 = s->Jaref[r] * mu, V0 = s->Jv[r] * mu, UU = 0, UV = 0, VV = 0, t0 = 0, t1 = 0, t2 = 0;
      for (int k = 0; k < dim; k++) { double Dk = o->efc_D[r + k], a0 = s->Jaref[r + k], av = s->Jv[r + k]; t0 += 0.5 * Dk * a0 * a0; t1 += Dk * a0 * av; t2 += 0.5 * Dk * av * av; }
      for (int k = 1; k < dim; k++) { double u = s->Jaref[r + k] * c->friction[k - 1], v = s->Jv[r + k] * c->friction[k - 1]; UU += u * u; UV += u * v; VV += v * v; }
      double N = U0 + a * V0, Tsqr = UU + a * (2 * UV + a * VV);
      if (Tsqr <= 0) { if (N < 0) { q0 += t0; q1 += t1; q2 += t2; } }
      else {
        double T = sqrt(Tsqr);
        if (N >= mu * T) { }
        else if (mu * N + T <= 0) { q0 += t0; q1 += t1; q2 += t2; }
        else {
          double Dm = o->efc_D[r] / fmax(MINVAL, mu * mu * (1 + mu * mu)), N1 = V0, T1 = (UV + a * VV) / T, T2 = VV / T - (UV + a * VV) * T1 / (T * T);
          double NmT = N - mu * T;
          p.cost += 0.5 * Dm * NmT * NmT; p.d1 += Dm * NmT * (N1 - mu * T1); p.d2 += Dm * ((N1 - mu * T1) * (N1 - mu * T1) + NmT * (-mu * T2));
        }
      }
      r += dim - 1;
    }
  }
  p.cost += a * a * q2 + a * q1 + q0; p.d1 += 2 * a * q2 + q1; p.d2 += 2 * q2;
  if (p.d2 < MINVAL) p.d2 = MINVAL;
  return p;
}
/* exact line search: Newton steps until the derivative changes sign, then safeguarded bracketing */
static double line_search(solver_t* s) {
  orc_t* o = s->o; int nv = o->nv;
  double snorm = 0; for (int i = 0; i < nv; i++) snorm += s->search[i] * s->search[i];
  snorm = sqrt(snorm);
  if (snorm < MINVAL) return 0;
  double gtol = o->tolerance * o->ls_tolerance * snorm / s->scale;
  mul_M(o, s->Mv, s->search);
  for (int r = 0; r < o->nefc; r++) s->Jv[r] = row_dot(o, r, s->search);
  double g1 = 0, g2 = 0, g0 = 0;
  for (int i = 0; i < nv; i++) { g1 += s->search[i] * (s->Ma[i] - o->qfrc_smooth[i]); g2 += 0.5 * s->search[i] * s->Mv[i]; g0 += 0.5 * (s->Ma[i] - o->qfrc_smooth[i]) * (o->qacc[i] - o->qacc_smooth[i]); }
  s->quadGauss[0] = g0; s->quadGauss[1] = g1; s->quadGauss[2] = g2;
  lspt_t p0 = ls_eval(s, 0), p1 = ls_eval(s, -p0.d1 / p0.d2), p2;
  if (p0.cost < p1.cost) p1 = p0;
  if (fabs(p1.d1) < gtol) return p1.alpha;
  int dir = p1.d1 < 0 ? 1 : -1, k = 0;
  p2 = p1;
  while (p1.d1 * dir <= -gtol && k < o->ls_iterations) {
    p2 = p1; p1 = ls_eval(s, p1.alpha - p1.d1 / p1.d2); k++;
    if (fabs(p1.d1) < gtol) return p1.alpha;
  }
  if (k >= o->ls_iterations) return p1.alpha;
  /* bracket: p2 on the descent side (d1*dir < 0), p1 beyond the minimum */
  lspt_t lo = p2, hi = p1, best = p1.cost < p2.cost ? p1 : p2;
  while (k < o->ls_iterations) {
    lspt_t* from = fabs(lo.d1) < fabs(hi.d1) ? &lo : &hi;
    double a = from->alpha - from->d1 / from->d2, amin = fmin(lo.alpha, hi.alpha), amax = fmax(lo.alpha, hi.alpha);
    if (!(a > amin && a < amax)) a = 0.5 * (lo.alpha + hi.alpha);
    lspt_t pm = ls_eval(s, a); k++;
    if (pm.cost < best.cost) best = pm;
    if (fabs(pm.d1) < gtol) return pm.alpha;
    if (pm.d1 * dir < 0) lo = pm; else hi = pm;
    if (fabs(hi.alpha - lo.alpha) < 1e-14 * (1 + fabs(hi.alpha))) break;
  }
  return best.alpha;
}
static double total_cost_at(orc_t* o, const double* qacc, double* jar, double* Ma) {
  int nv = o->nv;
  for (int r = 0; r < o->nefc; r++) jar[r] = row_dot(o, r, qacc) - o->efc_aref[r];
  mul_M(o, Ma, qacc);
  double cost = constraint_update(o, jar, NULL);
  for (int i = 0; i < nv; i++) cost += 0.5 * (Ma[i] - o->qfrc_smooth[i]) * (qacc[i] - o->qacc_smooth[i]);
  return cost;
}

/* ------------------------------------------------------------------ PGS solver (App. B item 14): the DUAL of the problem Newton solves.
 * minimise  1/2 f'(A + R)f + f'b   over  f in Omega,   A = J M^-1 J',  b = J qacc_smooth - aref,
 * Omega: equality rows free, limit rows f >= 0, elliptic contacts f_n >= 0, sum_k (f_tk / friction_k)^2 <= f_n^2 (the TRUE friction
 * cone; the primal solver works with the R-scaled cone and contact.mu instead).  Block Gauss-Seidel in MuJoCo's order: scalar rows
 * exactly; a contact by (1) a move along the ray through its current force, (2) the normal alone when the force sits at the apex,
 * else the friction components with the normal held fixed, a 2-variable QCQP (mju_QCQP2).  A second, independent minimiser of the
 * same convex problem: agreement with Newton checks R, D, contact.mu and the cone zones jointly (SURVEY 8c fixture 3). */
static void qcqp2(const double A[4], const double b[2], const double d[2], double r, double x[2]) {
  /* min 1/2 x'Ax + x'b  s.t. (x0/d0)^2 + (x1/d1)^2 <= r^2 ; scale y = x/d */
  double a00 = A[0] * d[0] * d[0], a01 = A[1] * d[0] * d[1], a11 = A[3] * d[1] * d[1], b0 = b[0] * d[0], b1 = b[1] * d[1];
  double la = 0, y0 = 0, y1 = 0;
  for (int it = 0; it < 40; it++) {
    double m00 = a00 + la, m11 = a11 + la, det = m00 * m11 - a01 * a01;
    if (det < 1e-300) break;
    y0 = -(m11 * b0 - a01 * b1) / det; y1 = -(-a01 * b0 + m00 * b1) / det;
    double val = y0 * y0 + y1 * y1 - r * r;
    if (val < 1e-14 * (1 + r * r)) break;                       /* inside (la = 0) or on the boundary */
    double z0 = (m11 * y0 - a01 * y1) / det, z1 = (-a01 * y0 + m00 * y1) / det;   /* (As + la I)^-1 y */
    double deriv = -2 * (y0 * z0 + y1 * z1);
    if (deriv > -1e-300) break;
    la -= val / deriv;
    if (la < 0) la = 0;
  }
  x[0] = y0 * d[0]; x[1] = y1 * d[1];
}
static void solve_pgs(orc_t* o) {
  int nv = o->nv, n = o->nefc;
  double* AR = (double*)calloc((size_t)n * n + (size_t)n * nv + 3 * (size_t)n, sizeof(double));
  double *X = AR + (size_t)n * n, *b = X + (size_t)n * nv, *f = b + n, *res = f + n;
  for (int r = 0; r < n; r++) {                                  /* X_r = M^-1 J_r' */
    memcpy(X + (size_t)r * nv, o->efc_J + (size_t)r * nv, 8 * nv);
    for (int i = 0; i < nv; i++) if (!o->dof_active[i]) X[(size_t)r * nv + i] = 0;
    solve_tree(o, o->qLD, X + (size_t)r * nv);
    for (int i = 0; i < nv; i++) if (!o->dof_active[i]) X[(size_t)r * nv + i] = 0;
  }
  for (int r = 0; r < n; r++) {
    for (int c = 0; c < n; c++) { double v = 0; const double* J = o->efc_J + (size_t)c * nv; for (int i = 0; i < nv; i++) v += J[i] * X[(size_t)r * nv + i]; AR[(size_t)r * n + c] = v; }
    AR[(size_t)r * n + r] += o->efc_R[r];
    double v = 0; const double* J = o->efc_J + (size_t)r * nv; for (int i = 0; i < nv; i++) v += J[i] * o->qacc_smooth[i];
    b[r] = v - o->efc_aref[r];
  }
  /* warm start: forces the primal map assigns to qacc_warmstart, if their dual cost beats f = 0 */
  {
    double* jar = res; double* Ma = (double*)calloc(nv, sizeof(double));
    total_cost_at(o, o->qacc_warmstart, jar, Ma); free(Ma);
    double c = 0;
    for (int r = 0; r < n; r++) { double v = 0; for (int k = 0; k < n; k++) v += AR[(size_t)r * n + k] * o->efc_force[k]; c += o->efc_force[r] * (0.5 * v + b[r]); }
    for (int r = 0; r < n; r++) f[r] = c < 0 ? o->efc_force[r] : 0;
  }
  double scale = 1 / (o->meaninertia * (nv > 1 ? nv : 1));
  int it = 0;
  for (; it < o->iterations; it++) {
    double improvement = 0;
    for (int r = 0; r < n; r++) {
      int dim = 1;
      if (o->efc_type[r] == ROW_CONTACT) dim = o->con[o->efc_id[r]].dim;
      double old[6], rs[6];
      for (int k = 0; k < dim; k++) { double v = b[r + k]; const double* a = AR + (size_t)(r + k) * n; for (int c = 0; c < n; c++) v += a[c] * f[c]; rs[k] = v; old[k] = f[r + k]; }
      if (dim == 1) {
        f[r] -= rs[0] / AR[(size_t)r * n + r];
        if (o->efc_type[r] != ROW_EQ && f[r] < 0) f[r] = 0;
      } else {
        contact_t* c = &o->con[o->efc_id[r]];
        double Ab[36];
        for (int i = 0; i < dim; i++) for (int j = 0; j < dim; j++) Ab[i * dim + j] = AR[(size_t)(r + i) * n + r + j];
        /* (1) ray update */
        if (f[r] > MINVAL) {
          double v1[6], den = 0, num = 0;
          for (int i = 0; i < dim; i++) { v1[i] = 0; for (int j = 0; j < dim; j++) v1[i] += Ab[i * dim + j] * old[j]; }
          for (int i = 0; i < dim; i++) { den += old[i] * v1[i]; num += old[i] * rs[i]; }
          if (den >= MINVAL) {
            double x = -num / den;
            if (f[r] + x * old[0] < 0) x = -1;                  /* do not leave the cone through the apex */
            for (int i = 0; i < dim; i++) f[r + i] += x * old[i];
          }
        }
        /* residual of the block with the rest fixed: bc = rs - Ab * old ; current block value v */
        double bc[6], v[6];
        for (int i = 0; i < dim; i++) { v[i] = f[r + i]; bc[i] = rs[i]; for (int j = 0; j < dim; j++) bc[i] -= Ab[i * dim + j] * old[j]; }
        if (v[0] < MINVAL) {
          /* (2) at the apex the gradient of the block cost is bc.  The apex is optimal iff -bc has no component inside the cone;
           * otherwise minimise along the projection d of -bc onto the cone (normal-only updates stall when the contact must start
           * sliding: d then lies on the cone boundary, not on its axis) */
          double mu1 = c->friction[0], an = -bc[0], bt[2] = {-bc[1], -bc[2]}, nb = sqrt(bt[0] * bt[0] + bt[1] * bt[1]), d[3] = {0, 0, 0};
          if (nb <= mu1 * an) { d[0] = an; d[1] = bt[0]; d[2] = bt[1]; }
          else if (mu1 * nb > -an) { d[0] = (an + mu1 * nb) / (1 + mu1 * mu1); d[1] = mu1 * d[0] * bt[0] / nb; d[2] = mu1 * d[0] * bt[1] / nb; }
          double den = 0, num = 0;
          for (int i = 0; i < 3; i++) { num += d[i] * bc[i]; for (int j = 0; j < 3; j++) den += d[i] * Ab[i * dim + j] * d[j]; }
          double x = den > MINVAL ? -num / den : 0;
          if (x < 0) x = 0;
          for (int i = 0; i < 3; i++) { v[i] = x * d[i]; f[r + i] = v[i]; }
        }
        if (v[0] >= MINVAL) {                                   /* (3) friction with the normal fixed (QCQP) */
          if (dim != 3) { fprintf(stderr, "orc: PGS handles condim 1 and 3 only\n"); }
          double A2[4] = {Ab[1 * dim + 1], Ab[1 * dim + 2], Ab[2 * dim + 1], Ab[2 * dim + 2]};
          double b2[2] = {bc[1] + Ab[1 * dim] * v[0], bc[2] + Ab[2 * dim] * v[0]}, d2[2] = {c->friction[0], c->friction[1]}, x2[2];
          qcqp2(A2, b2, d2, v[0], x2);
          f[r + 1] = x2[0]; f[r + 2] = x2[1];
        }
      }
      /* cost change of this block: dc = 1/2 d'Ab d + d'rs */
      double dc = 0;
      for (int i = 0; i < dim; i++) { double di = f[r + i] - old[i]; dc += di * rs[i]; for (int j = 0; j < dim; j++) dc += 0.5 * di * AR[(size_t)(r + i) * n + r + j] * (f[r + j] - old[j]); }
      if (dc > 1e-10) { for (int i = 0; i < dim; i++) f[r + i] = old[i]; dc = 0; }   /* never accept an increase */
      improvement -= dc;
      r += dim - 1;
    }
    o->solver_stats[0] = improvement * scale;
    if (improvement * scale < o->tolerance) { it++; break; }
  }
  o->solver_iter = it;
  memcpy(o->efc_force, f, 8 * n);
  memset(o->qfrc_constraint, 0, 8 * nv);
  for (int r = 0; r < n; r++) if (f[r] != 0) { const double* J = o->efc_J + (size_t)r * nv; for (int i = 0; i < nv; i++) o->qfrc_constraint[i] += J[i] * f[r]; }
  for (int i = 0; i < nv; i++) o->qacc[i] = o->dof_active[i] ? o->qfrc_constraint[i] : 0;
  solve_tree(o, o->qLD, o->qacc);
  for (int i = 0; i < nv; i++) o->qacc[i] = o->dof_active[i] ? o->qacc[i] + o->qacc_smooth[i] : o->qacc_smooth[i];
  free(AR);
}
static void fwd_constraint(orc_t* o) {
  int nv = o->nv; o->solver_iter = 0;
  if (o->nefc == 0) { memcpy(o->qacc, o->qacc_smooth, 8 * nv); memset(o->qfrc_constraint, 0, 8 * nv); return; }
  if (o->solver == 1) { solve_pgs(o); return; }
  double* buf = o->solbuf; memset(buf, 0, sizeof(double) * ((size_t)2 * ORC_MAXEFC + 6 * MAXNV));
  solver_t s; s.o = o; s.Jaref = buf; s.Jv = buf + ORC_MAXEFC; s.Ma = s.Jv + ORC_MAXEFC; s.Mv = s.Ma + MAXNV; s.search = s.Mv + MAXNV;
  s.grad = s.search + MAXNV; s.Mgrad = s.grad + MAXNV; s.H = s.Mgrad + MAXNV + MAXNV;
  s.scale = 1 / (o->meaninertia * (nv > 1 ? nv : 1));
  /* warm start: the cheaper of qacc_warmstart and qacc_smooth */
  double cw = total_cost_at(o, o->qacc_warmstart, s.Jaref, s.Ma), cs = total_cost_at(o, o->qacc_smooth, s.Jaref, s.Ma);
  memcpy(o->qacc, cw < cs ? o->qacc_warmstart : o->qacc_smooth, 8 * nv);
  total_cost_at(o, o->qacc, s.Jaref, s.Ma);
  solver_update(&s, o->qacc);
  for (int i = 0; i < nv; i++) s.search[i] = -s.Mgrad[i];
  int it = 0;
  for (; it < o->iterations; it++) {
    double alpha = line_search(&s);
    if (alpha == 0) break;
    for (int i = 0; i < nv; i++) { o->qacc[i] += alpha * s.search[i]; s.Ma[i] += alpha * s.Mv[i]; }
    for (int r = 0; r < o->nefc; r++) s.Jaref[r] += alpha * s.Jv[r];
    double old = s.cost;
    solver_update(&s, o->qacc);
    double gn = 0; for (int i = 0; i < nv; i++) gn += s.grad[i] * s.grad[i];
    double improvement = s.scale * (old - s.cost), gradient = s.scale * sqrt(gn);
    o->solver_stats[0] = improvement; o->solver_stats[1] = gradient;
    if (improvement < o->tolerance || gradient < o->tolerance) { it++; break; }
    for (int i = 0; i < nv; i++) s.search[i] = -s.Mgrad[i];
  }
  o->solver_iter = it; o->solver_stats[2] = s.cost;
}

/* ------------------------------------------------------------------ pipeline */
static void fwd_position(orc_t* o) {
  kinematics(o); com_pos(o); crb(o);
  memcpy(o->qLD, o->qM, 8 * o->nv * o->nv); factor_tree(o, o->qLD);
  collision(o); make_constraint(o);
}
static void fwd_velocity(orc_t* o) { com_vel(o); passive(o); reference_constraint(o); rne_bias(o); }
static void fwd_acceleration(orc_t* o) {
  for (int i = 0; i < o->nv; i++) {
    o->qfrc_smooth[i] = o->dof_active[i] ? o->qfrc_passive[i] - o->qfrc_bias[i] + o->qfrc_actuator[i] : 0;
    o->qacc_smooth[i] = o->qfrc_smooth[i];
  }
  solve_tree(o, o->qLD, o->qacc_smooth);
}
void orc_step1(orc_t* o) { fwd_position(o); fwd_velocity(o); }
static void integrate(orc_t* o) {
  /* mj_implicit (implicitfast) / mj_Euler + mj_advance (App. B item 15) */
  int nv = o->nv; double h = o->timestep, qacc[MAXNV];
  const double *damp = FD(o, B2PM_F_DOF_DAMPING), *bias = FD(o, B2PM_F_ACT_BIAS), *tc = FD(o, B2PM_F_ACT_TRNCOEF);
  const int *tn = FI(o, B2PM_F_ACT_TRNNUM), *td = FI(o, B2PM_F_ACT_TRNDOF), *dpar = FI(o, B2PM_F_DOF_PARENTID);
  memcpy(o->qH, o->qM, 8 * nv * nv);
  for (int i = 0; i < nv; i++) o->qH[i * nv + i] += h * damp[i];
  if (o->integrator == B2P_INT_IMPLICITFAST) {
    /* actuator velocity derivative moment' * bias_vel * moment, restricted to M's (ancestor) sparsity */
    for (int a = 0; a < o->nu; a++) if (bias[3 * a + 2] != 0) for (int x = 0; x < tn[a]; x++) for (int y = 0; y < tn[a]; y++) {
      int i = td[2 * a + x], j = td[2 * a + y], anc = 0;
      for (int k = i; k >= 0; k = dpar[k]) if (k == j) anc = 1;
      for (int k = j; k >= 0; k = dpar[k]) if (k == i) anc = 1;
      if (anc) o->qH[i * nv + j] -= h * tc[2 * a + x] * tc[2 * a + y] * bias[3 * a + 2];
    }
  }
  factor_tree(o, o->qH);
  for (int i = 0; i < nv; i++) qacc[i] = o->dof_active[i] ? o->qfrc_smooth[i] + o->qfrc_constraint[i] : 0;
  solve_tree(o, o->qH, qacc);
  for (int i = 0; i < nv; i++) o->qvel[i] += h * qacc[i];
  const int *jt = FI(o, B2PM_F_JNT_TYPE), *jq = FI(o, B2PM_F_JNT_QPOSADR), *jd = FI(o, B2PM_F_JNT_DOFADR);
  for (int j = 0; j < o->njnt; j++) {
    int qa = jq[j], d = jd[j];
    if (jt[j] == B2P_JNT_HINGE) o->qpos[qa] += h * o->qvel[d];
    else {
      for (int k = 0; k < 3; k++) o->qpos[qa + k] += h * o->qvel[d + k];
      double w[3] = {o->qvel[d + 3], o->qvel[d + 4], o->qvel[d + 5]}, ang = h * normalize3(w);
      double qr[4] = {cos(0.5 * ang), sin(0.5 * ang) * w[0], sin(0.5 * ang) * w[1], sin(0.5 * ang) * w[2]};
      mulquat(o->qpos + qa + 3, o->qpos + qa + 3, qr); normquat(o->qpos + qa + 3);
    }
  }
  memcpy(o->qacc_warmstart, o->qacc, 8 * nv);
  o->time += h;
}
static void step2_noint(orc_t* o) { actuation(o); fwd_acceleration(o); fwd_constraint(o); }
void orc_step2(orc_t* o) { step2_noint(o); integrate(o); }
void orc_step(orc_t* o) { orc_step1(o); orc_step2(o); }
void orc_forward(orc_t* o) { orc_step1(o); step2_noint(o); }
void orc_step_frozen_arm(orc_t* o) {
  /* dm_control JointStaticIsolator around physics.step() (prop_initializer.py:245-250, App. B item 18) */
  int na = o->hdr[B2PM_C_NARM] + o->hdr[B2PM_C_NGRIP]; double q[MAXNV], v[MAXNV];
  memcpy(q, o->qpos, 8 * na); memcpy(v, o->qvel, 8 * na);
  orc_step(o);
  memcpy(o->qpos, q, 8 * na); memcpy(o->qvel, v, 8 * na);
}
int orc_contacts(orc_t* o, double* dist, double* pos, double* frame, int* geom, int* excluded) {
  for (int i = 0; i < o->ncon; i++) {
    if (dist) dist[i] = o->con[i].dist;
    if (pos) memcpy(pos + 3 * i, o->con[i].pos, 24);
    if (frame) memcpy(frame + 9 * i, o->con[i].frame, 72);
    if (geom) { geom[2 * i] = o->con[i].geom1; geom[2 * i + 1] = o->con[i].geom2; }
    if (excluded) excluded[i] = o->con[i].exclude;
  }
  return o->ncon;
}

/* ------------------------------------------------------------------ OSC (mujoco_controllers/osc.py) */
void orc_osc_config(orc_t* o, const double g[6], const double q0[7], double pt, double ot) {
  memcpy(o->osc_gains, g, 48); memcpy(o->osc_null, q0, 56); o->osc_pos_thresh = pt; o->osc_ori_thresh = ot;
}
static void inv_spd(double* A, int n, double* Ainv) { /* via Cholesky; A destroyed */
  cholesky(A, n);
  for (int c = 0; c < n; c++) { double e[8] = {0}; e[c] = 1; chol_solve(A, n, e); for (int r = 0; r < n; r++) Ainv[r * n + c] = e[r]; }
}
static void jacobi_eig(double* A, int n, double* V) { /* symmetric A -> diag in place, V columns = eigenvectors */
  for (int i = 0; i < n; i++) for (int j = 0; j < n; j++) V[i * n + j] = i == j;
  for (int sweep = 0; sweep < 60; sweep++) {
    double off = 0; for (int i = 0; i < n; i++) for (int j = i + 1; j < n; j++) off += A[i * n + j] * A[i * n + j];
    if (off < 1e-300) break;
    for (int p = 0; p < n; p++) for (int q = p + 1; q < n; q++) {
      if (fabs(A[p * n + q]) < 1e-300) continue;
      double th = (A[q * n + q] - A[p * n + p]) / (2 * A[p * n + q]), t = (th >= 0 ? 1 : -1) / (fabs(th) + sqrt(th * th + 1)), c = 1 / sqrt(t * t + 1), sn = t * c;
      for (int k = 0; k < n; k++) { double a = A[k * n + p], b = A[k * n + q]; A[k * n + p] = c * a - sn * b; A[k * n + q] = sn * a + c * b; }
      for (int k = 0; k < n; k++) { double a = A[p * n + k], b = A[q * n + k]; A[p * n + k] = c * a - sn * b; A[q * n + k] = sn * a + c * b; }
      for (int k = 0; k < n; k++) { double a = V[k * n + p], b = V[k * n + q]; V[k * n + p] = c * a - sn * b; V[k * n + q] = sn * a + c * b; }
    }
  }
}
static void osc_orientation_error(const double* q, const double* qd, double* e) { /* osc.py:132-150 */
  double qc[4] = {q[0], -q[1], -q[2], -q[3]}, qe[4]; normquat(qc);
  mulquat(qe, qd, qc);
  double sg = qe[0] > 0 ? 1 : (qe[0] < 0 ? -1 : 0);
  e[0] = qe[1] * sg; e[1] = qe[2] * sg; e[2] = qe[3] * sg;
}
static void osc_eef_quat(orc_t* o, double* q) { mat2quat(q, o->site_xmat + 9 * o->hdr[B2PM_C_SITE_ATTACHMENT]); }
void orc_osc_torque(orc_t* o, const double tg[13], double tau[7]) {
  int nv = o->nv, site = o->hdr[B2PM_C_SITE_ATTACHMENT], body = FI(o, B2PM_F_SITE_BODYID)[site];
  double* jp = (double*)malloc(sizeof(double) * 6 * nv); double* jr = jp + 3 * nv;
  jac_point(o, body, o->site_xpos + 3 * site, jp, jr);                         /* osc.py:117-130 */
  double J[42], Ma[49], Minv[49], MiJt[42], Lxi[36], Lx[36];
  for (int k = 0; k < 3; k++) for (int i = 0; i < 7; i++) { J[k * 7 + i] = jp[k * nv + i]; J[(3 + k) * 7 + i] = jr[k * nv + i]; }
  free(jp);
  for (int i = 0; i < 7; i++) for (int j = 0; j < 7; j++) Ma[i * 7 + j] = o->qM[i * nv + j];        /* osc.py:99-105 */
  inv_spd(Ma, 7, Minv);
  for (int i = 0; i < 7; i++) for (int r = 0; r < 6; r++) { double s = 0; for (int k = 0; k < 7; k++) s += Minv[i * 7 + k] * J[r * 7 + k]; MiJt[i * 6 + r] = s; }
  for (int r = 0; r < 6; r++) for (int c = 0; c < 6; c++) { double s = 0; for (int k = 0; k < 7; k++) s += J[r * 7 + k] * MiJt[k * 6 + c]; Lxi[r * 6 + c] = s; }
  double L[36]; memcpy(L, Lxi, sizeof L);
  double det = 1; int ok = cholesky(L, 6) == 0;
  if (ok) for (int i = 0; i < 6; i++) det *= L[i * 6 + i] * L[i * 6 + i]; else det = 0;
  if (fabs(det) >= 1e-2) { memcpy(L, Lxi, sizeof L); inv_spd(L, 6, Lx); }       /* osc.py:109-112 */
  else {                                                                       /* osc.py:113-115 pinv(rcond=1e-2) */
    double A[36], V[36]; memcpy(A, Lxi, sizeof A); jacobi_eig(A, 6, V);
    double smax = 0; for (int i = 0; i < 6; i++) if (fabs(A[i * 6 + i]) > smax) smax = fabs(A[i * 6 + i]);
    for (int r = 0; r < 6; r++) for (int c = 0; c < 6; c++) { double s = 0; for (int k = 0; k < 6; k++) { double ev = A[k * 6 + k]; if (fabs(ev) > 1e-2 * smax) s += V[r * 6 + k] * V[c * 6 + k] / ev; } Lx[r * 6 + c] = s; }
  }
  double x[3], q[4], dx[6] = {0};
  memcpy(x, o->site_xpos + 3 * site, 24); osc_eef_quat(o, q);
  for (int r = 0; r < 6; r++) for (int i = 0; i < 7; i++) dx[r] += J[r * 7 + i] * o->qvel[i];
  const double* g = o->osc_gains; double pd[6], eo[3];
  for (int k = 0; k < 3; k++) { double e = fmin(0.05, fmax(-0.05, tg[k] - x[k])); pd[k] = g[0] * e + g[1] * (tg[3 + k] - dx[k]); }   /* osc.py:168-172 */
  double qd[4] = {tg[6], tg[7], tg[8], tg[9]}; normquat(qd);
  osc_orientation_error(q, qd, eo);
  for (int k = 0; k < 3; k++) pd[3 + k] = g[2] * eo[k] + g[3] * (tg[10 + k] - dx[3 + k]);
  double Lpd[6];
  for (int r = 0; r < 6; r++) { double s = 0; for (int c = 0; c < 6; c++) s += Lx[r * 6 + c] * pd[c]; Lpd[r] = s; }
  for (int i = 0; i < 7; i++) { double s = 0; for (int r = 0; r < 6; r++) s += J[r * 7 + i] * Lpd[r]; tau[i] = s; }               /* osc.py:206-207 */
  double ne[7], Nb[42]; /* null_jacobian = Minv J' Lx  (7x6)  osc.py:218 */
  for (int i = 0; i < 7; i++) ne[i] = g[4] * (o->osc_null[i] - o->qpos[i]) + g[5] * (0 - o->qvel[i]);
  for (int i = 0; i < 7; i++) for (int c = 0; c < 6; c++) { double s = 0; for (int r = 0; r < 6; r++) s += MiJt[i * 6 + r] * Lx[r * 6 + c]; Nb[i * 6 + c] = s; }
  for (int i = 0; i < 7; i++) {
    double s = ne[i];
    for (int j = 0; j < 7; j++) { double jn = 0; for (int c = 0; c < 6; c++) jn += J[c * 7 + i] * Nb[j * 6 + c]; s -= jn * ne[j]; }   /* (I - J' Nb') ne */
    tau[i] += s + o->qfrc_bias[i];                                                                                             /* osc.py:219-222 */
  }
  /* osc.py:224-236: tau @ pinv(actuator_moment)[arm,arm] == tau for unit-gear joint motors */
  memcpy(o->osc_tau, tau, 56);
}
int orc_osc_converged(orc_t* o, const double tg[13]) { /* osc.py:240-247 */
  int site = o->hdr[B2PM_C_SITE_ATTACHMENT]; double q[4], e[3], d[3], qd[4] = {tg[6], tg[7], tg[8], tg[9]};
  normquat(qd); osc_eef_quat(o, q); osc_orientation_error(q, qd, e);
  for (int k = 0; k < 3; k++) d[k] = o->site_xpos[3 * site + k] - tg[k];
  double oe = fmax(fabs(e[0]), fmax(fabs(e[1]), fabs(e[2])));
  return norm3(d) < o->osc_pos_thresh && oe < o->osc_ori_thresh;
}
int orc_run_controller(orc_t* o, const double tg[13], int grip_max, int n_ticks, int control_steps) {
  /* models/robot_arm.py:61-94 with dm_control's legacy step (mj_step2 ; mj_step1): derived quantities are
   * consistent with the current state whenever the controller reads them */
  int converged = 0;
  orc_step1(o);
  for (int t = 0; t < n_ticks; t++) {
    double tau[7]; orc_osc_torque(o, tg, tau);
    memcpy(o->ctrl, tau, 56); o->ctrl[7] = grip_max ? 255.0 : 0.0;          /* min_max.py:14-19 */
    for (int k = 0; k < control_steps; k++) { orc_step2(o); orc_step1(o); }
    if (orc_osc_converged(o, tg)) converged = 1;
  }
  return converged;
}
