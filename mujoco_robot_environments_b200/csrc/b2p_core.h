// B200-native batched physics + OSC for the rearrangement model class: per-env algorithms.
//
// One CTA steps one scene ("block-per-env"): all scratch of a scene lives in shared memory
// (Scratch<T>), model constants are staged once per CTA (DevModel<T>), parallel loops are written
// as `for (i = tid; i < n; i += nt)` separated by block barriers.  The same templates compile
// for the host with nt == 1 (barriers become no-ops): that build exists ONLY so the CPU test-suite
// can check the kernel logic against the oracle without a GPU (tests/_hostemu); the product path
// is the CUDA build in b2p_api.cu and has no CPU fallback.
//
// Replaces, for this model class, MuJoCo 3.2.7's mj_step1/mj_step2 as driven by the reference's
// RobotArm.run_controller (mujoco_robot_environments/models/robot_arm.py:61-94) and the OSC maths of
// mujoco_controllers/osc.py:99-247.  Stage semantics: SURVEY.md Appendix B.
#pragma once
#include "b2p_model.h"

#include "b2p_kin.h"
#include "b2p_collide.h"
#include "b2p_rows.h"
#include "b2p_solver.h"
#include "b2p_osc.h"
#include "b2p_pgs.h"

namespace b2p {

// ------------------------------------------------------------------------------------------ pipeline
template <typename T, bool C>
B2P_HD void prop_params(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c) {
  // mass, principal inertia, invweight0 and bounding half extents per collider class (include/b2p_prop_types.h: b2p_prop_inertia)
  for (int p = c.tid; p < m.nprop; p += c.nt) {
    const T* z = s.psize[p]; const int tp = s.ptype[p];
    T mass, I[3], hf[3];
    if (C && tp == B2P_GEOM_SPHERE) {
      const T r = z[0]; mass = m.pclass_mass[0] > 0 ? m.pclass_mass[0] : T(1000.0 * 4.0 / 3.0 * 3.14159265358979323846) * r * r * r;
      I[0] = I[1] = I[2] = T(0.4) * mass * r * r; hf[0] = hf[1] = hf[2] = r;
    } else if (C && tp == B2P_GEOM_CYLINDER) {
      const T r = z[0], h = z[1]; mass = m.pclass_mass[1] > 0 ? m.pclass_mass[1] : T(1000.0 * 3.14159265358979323846) * r * r * 2 * h;
      I[0] = I[1] = mass * (T(0.25) * r * r + h * h / 3); I[2] = T(0.5) * mass * r * r; hf[0] = hf[1] = r; hf[2] = h;
    } else {
      mass = m.prop_mass[p];
      I[0] = mass / 3 * (z[1] * z[1] + z[2] * z[2]); I[1] = mass / 3 * (z[0] * z[0] + z[2] * z[2]); I[2] = mass / 3 * (z[0] * z[0] + z[1] * z[1]);
      hf[0] = z[0]; hf[1] = z[1]; hf[2] = z[2];
    }
    for (int k = 0; k < 3; k++) { s.pinertia[p][k] = I[k]; s.phalf[p][k] = hf[k]; }
    s.pmass[p] = mass;
    s.pinvw[p][0] = 1 / mass; s.pinvw[p][1] = (1 / I[0] + 1 / I[1] + 1 / I[2]) / 3;
  }
  bsync(c);
}
template <typename T, bool C>
B2P_HDN void step1(const DevModel<T>& m, const Globals<T>& g, Scratch<T, C>& s, const Ctx& c, bool full_contacts, bool active) {
  PROF_T0
  csync(c, LK_FK);
  PROF_ADD(PF_BARRIER)
  if (active) { forward_sweep(m, s, c); PROF_ADD(PF_FK) backward_sweep(m, s, c); dynamics_terms(m, s, c); PROF_ADD(PF_DYN) }
  csync(c, LK_COLLIDE);
  PROF_ADD(PF_BARRIER)
  collide(m, g, s, c, full_contacts, active);
  PROF_ADD(PF_COLLIDE)
  csync(c, LK_ROWS);
  PROF_ADD(PF_BARRIER)
  if (active) make_rows(m, g, s, c);
  PROF_ADD(PF_ROWS)
}
template <typename T, bool kPgs = false, bool C>
B2P_HDN void step2(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c, bool do_integrate, bool freeze_tree, bool active) {
  PROF_T0
  csync(c, LK_STEP2);
  PROF_ADD(PF_BARRIER)
  if (active) {
    PROF_COUNT(PF_NSTEP, 1)
    actuation_and_smooth(m, s, c);
    NANCHK(1, s.qacc_smooth, m.nv)
  }
  PROF_ADD(PF_SMOOTH)
  csync(c, LK_SOLVE);
  if constexpr (kPgs) solve_pgs(m, s, c, active); else solve_constraints(m, s, c, active);
  PROF_RESET
  if (active) {
    // mj_checkAcc / checkVel / checkPos equivalents -> per-env status bits
    unsigned bad = 0;
    for (int d = c.tid; d < m.nv; d += c.nt) { if (!(fabs(s.qacc[d]) < T(1e10))) bad |= FLAG_BADQACC; if (!(fabs(s.qvel[d]) < T(1e10))) bad |= FLAG_BADQVEL; }
    for (int d = c.tid; d < m.nq; d += c.nt) if (!(fabs(s.qpos[d]) < T(1e10))) bad |= FLAG_BADQPOS;
    bad = block_or(c, bad);
    if (c.tid == 0 && bad) s.flags |= (int)bad;
  }
  PROF_ADD(PF_MISC)
  csync(c, LK_INTEGRATE);
  PROF_ADD(PF_BARRIER)
  if (active) { if (do_integrate) integrate(m, s, c, freeze_tree); else bsync(c); }
  PROF_ADD(PF_INTEGRATE)
}

// Derived-output record, written group by group (b2p_set_out_mask; B2P_OUT_* of include/b2p.h).  The "pos" groups (site / geom
// poses, contacts, mass matrix, bias) read step1 buffers that share storage with the solver's (Scratch unions), so they must be
// written after a step1 and before the next step2; the "dyn" groups (qacc, efc_force, solver statistics) persist.
template <typename T, bool C>
B2P_HDN void write_out(const DevModel<T>& m, const Globals<T>& g, const Scratch<T, C>& s, const Ctx& c, T* o, int parts) {
  typedef OutLayout L;
  if (parts & OUT_SITE) {
    for (int i = c.tid; i < 6; i += c.nt) o[L::SITE_XPOS + i] = s.site_xpos[i / 3][i % 3];
    for (int i = c.tid; i < 18; i += c.nt) o[L::SITE_XMAT + i] = s.site_xmat[i / 9][i % 9];
    if (c.tid == 0) { o[L::NCON] = T(s.ncon); o[L::NEFC] = T(s.nefc); }
  }
  if (parts & OUT_BIAS) for (int i = c.tid; i < MAXNV; i += c.nt) o[L::QFRC_BIAS + i] = i < m.nv ? s.qfrc_bias[i] : T(0);
  if (parts & OUT_CONTACT) {
    for (int k = c.tid; k < MAXCON; k += c.nt) {
      T* q = o + L::CONTACT + k * L::CONTACT_W;
      if (k < s.ncon) {
        const int p = s.c_pair[k];
        q[0] = s.c_dist[k]; for (int i = 0; i < 3; i++) { q[1 + i] = s.c_pos[k][i]; q[4 + i] = s.c_frame[k][i]; }
        q[7] = T(g.pair_g1[p]); q[8] = T(g.pair_g2[p]); q[9] = T(s.c_dist[k] < s.c_incl[k] ? 0 : 1);
      } else for (int i = 0; i < L::CONTACT_W; i++) q[i] = 0;
    }
  }
  if (parts & OUT_MARM) for (int i = c.tid; i < MAXTD * MAXTD; i += c.nt) { const int r = i / MAXTD, q = i % MAXTD; o[L::MARM + i] = s.Mt[r >= q ? tri_idx(r, q) : tri_idx(q, r)]; }
  if (parts & OUT_GEOM) {
    for (int i = c.tid; i < MAXG * 3; i += c.nt) o[L::GPOS + i] = i / 3 < m.ngeom ? s.gpos[i / 3][i % 3] : T(0);
    for (int i = c.tid; i < MAXG * 9; i += c.nt) o[L::GMAT + i] = i / 9 < m.ngeom ? s.gmat[i / 9][i % 9] : T(0);
  }
  if (parts & OUT_DYN) {
    for (int i = c.tid; i < MAXNV; i += c.nt) { o[L::QACC + i] = i < m.nv ? s.qacc[i] : T(0); o[L::QACC_SMOOTH + i] = i < m.nv ? s.qacc_smooth[i] : T(0); }
    for (int i = c.tid; i < MAXU; i += c.nt) o[L::CTRL + i] = s.ctrl[i];
    if (c.tid == 0) o[L::NITER] = T(s.niter);
  }
  if (parts & OUT_EFC) for (int i = c.tid; i < MAXEFC_OUT; i += c.nt) o[L::EFC_FORCE + i] = i < s.nefc ? s.force[i] : T(0);
}

// One scene, whole launch: load the state record, run the requested mode, store the record.  `active` = this warp holds
// a scene; an idle warp of a lockstep CTA still walks the same stage barriers.
// kVar: 0 = the step kernel, 1 = the placement search (MODE_PLACE), 2 = the step kernel with the PGS solver.  Each is its own
// instantiation, so the default step kernel's code (and its layout in the instruction cache) is untouched by the other two.
enum { VAR_STEP = 0, VAR_PLACE = 1, VAR_PGS = 2 };
template <typename T, int kVar = VAR_STEP, bool C>
B2P_HD void run_env(const DevModel<T>& m, const Globals<T>& g, Scratch<T, C>& s, const StepArgs<T>& a, int env, const Ctx& c, bool state_preloaded, bool active = true) {
  constexpr bool kPlace = kVar == VAR_PLACE, kPgs = kVar == VAR_PGS;
  const int nv = m.nv;
  T* rec = a.state + (size_t)(active ? env : 0) * a.state_stride;
  if (active) {
    if (!state_preloaded) {   // the CUDA kernel stages the record with one TMA bulk copy instead
      T* head = s.qpos;
      for (int i = c.tid; i < REC_SIZE; i += c.nt) head[i] = rec[i];
    }
    for (int i = c.tid; i < MAXP * 3; i += c.nt) s.psize[i / 3][i % 3] = a.prop_size ? a.prop_size[(size_t)env * MAXP * 3 + i] : m.geom_size[0][0];
    for (int i = c.tid; i < MAXP; i += c.nt) s.pactive[i] = i < m.nprop ? (a.prop_active ? a.prop_active[(size_t)env * MAXP + i] : 1) : 0;
    for (int i = c.tid; i < MAXP; i += c.nt) { const int tp = (C && a.prop_type && i < m.nprop) ? a.prop_type[(size_t)env * MAXP + i] : B2P_GEOM_BOX; s.ptype[i] = (unsigned char)((tp == B2P_GEOM_SPHERE || tp == B2P_GEOM_CYLINDER) ? tp : B2P_GEOM_BOX); }
    for (int i = c.tid; i < MAXU; i += c.nt) s.ctrl[i] = (a.ctrl && i < m.nu) ? a.ctrl[(size_t)env * m.nu + i] : T(0);
    if (c.tid == 0) { s.flags = 0; s.niter = 0; s.ncon = 0; s.nefc = 0; s.sp.nblk = 0; s.sep_next = 0; for (int k = 0; k < SEP_N; k++) s.sep_pair[k] = -1; }   // sp.nblk = 0: no block structure built yet (make_rows)
    bsync(c);
    if (!a.prop_size) {   // default: sizes from the model's prop geoms
      for (int g0 = c.tid; g0 < m.ngeom; g0 += c.nt) if (m.geom_prop[g0] >= 0) for (int k = 0; k < 3; k++) s.psize[m.geom_prop[g0]][k] = m.geom_size[g0][k];
      bsync(c);
    }
    for (int g0 = c.tid; g0 < MAXG; g0 += c.nt) s.gtype[g0] = (unsigned char)(g0 < m.ngeom ? (m.geom_prop[g0] >= 0 ? s.ptype[m.geom_prop[g0]] : m.geom_type[g0]) : 0);
    if (c.tid == 0) { int any = 0; for (int p = 0; p < m.nprop; p++) any |= s.ptype[p] != B2P_GEOM_BOX; s.reclassed = any; }
    prop_params(m, s, c);
  }
  if constexpr (kPlace) {
    // SortColours prop_place (rearrangement.py:597-665) as one launch: draw a position in the scene's box, put the prop there
    // (gripper-down orientation), run kinematics + the full collision pass, accept when no contact of the prop with anything
    // but the table is closer than place_dist.  The live state is not written back (the reference works on a deepcopy).
    int gid = -1, qa = m.ntd; T lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
    if (active) {
      const int pid = a.place_prop[env]; qa = m.ntd + 7 * pid;
      for (int g0 = 0; g0 < m.ngeom; g0++) if (m.geom_prop[g0] == pid) gid = g0;
      for (int k = 0; k < 3; k++) { lo[k] = a.place_box[(size_t)env * 6 + k]; hi[k] = a.place_box[(size_t)env * 6 + 3 + k]; }
    }
    bool fin = !active, found = false; int used = 0;
    for (int k = 0; k < a.n_steps; k++) {
      if (csync_and(c, fin)) break;
      if (!fin) {
        if (c.tid == 0) {
          T* q = s.qpos + qa;
          for (int d = 0; d < 3; d++) q[d] = lo[d] + (hi[d] - lo[d]) * (T)uniform01(a.place_seed, (unsigned)env, (unsigned)k, (unsigned)d);
          if (a.place_yaw) {   // PropPlacer (rearrangement.py:189-197): yaw = pi * U about z, quaternion (cos, 0, 0, sin)
            const T ang = T(3.14159265358979323846) * (T)uniform01(a.place_seed, (unsigned)env, (unsigned)k, 3u);
            q[3] = cos(T(0.5) * ang); q[4] = 0; q[5] = 0; q[6] = sin(T(0.5) * ang);
          } else { q[3] = 0; q[4] = 0; q[5] = 1; q[6] = 0; }
        }
        bsync(c);
      }
      csync(c, LK_FK);
      if (!fin) forward_sweep(m, s, c);
      csync(c, LK_COLLIDE);
      collide(m, g, s, c, true, !fin);
      if (!fin) {
        bsync(c);
        unsigned hit = 0;
        for (int i = c.tid; i < s.ncon; i += c.nt) {
          const int p = s.c_pair[i], g1 = g.pair_g1[p], g2 = g.pair_g2[p];
          if ((g1 == gid || g2 == gid) && g1 != a.place_table && g2 != a.place_table && s.c_dist[i] <= a.place_dist) hit = 1;
        }
        hit = block_or(c, hit);
        used = k + 1;
        if (!hit) { found = true; fin = true; }
      }
    }
    if (active && c.tid == 0 && a.place_out) {
      T* o = a.place_out + (size_t)env * 4;
      for (int d = 0; d < 3; d++) o[d] = s.qpos[qa + d];
      o[3] = found ? T(used) : T(-used - 1);
      if (a.place_quat) for (int d = 0; d < 4; d++) a.place_quat[(size_t)env * 4 + d] = s.qpos[qa + 3 + d];
    }
    return;
  }
  const bool full = a.full_contacts != 0;
  int work = 0;   // scheduling hint for the next launch (see StepArgs::work)
  if (a.mode == MODE_FORWARD) {
    step1(m, g, s, c, full, active);
    if (active && a.out) { bsync(c); write_out(m, g, s, c, a.out + (size_t)env * a.out_stride, OUT_POS); bsync(c); }   // physics.forward(): every group
    step2<T, kPgs>(m, s, c, false, false, active);
  } else if (a.mode == MODE_PHYSICS) {
    for (int k = 0; k < a.n_steps; k++) { step1(m, g, s, c, full, active); step2<T, kPgs>(m, s, c, true, false, active); if (active) work += s.nefc * (2 + s.niter); }
    step1(m, g, s, c, full, active);   // leave derived quantities consistent with the final state (dm_control legacy step order)
  } else if (a.mode == MODE_OSC) {
    // RobotArm.run_controller (robot_arm.py:61-94): per tick OSC torque + gripper command, control_steps x physics.step()
    T tg[13];
    for (int i = 0; i < 13; i++) tg[i] = active ? a.target[(size_t)env * 13 + i] : T(0);
    const bool grip = (active && a.grip) ? a.grip[env] != 0 : false;
    T gn[6];   // controller_gains: the handle's (osc.yaml) or this scene's row (the tuner evaluates one candidate per scene)
    for (int i = 0; i < 6; i++) gn[i] = (active && a.gains) ? a.gains[(size_t)env * 6 + i] : m.osc_gains[i];
    bool conv = false;
    step1(m, g, s, c, full, active);
    if (a.n_steps == 0 && active) {   // OSC.compute_control_output() / is_converged() for the current state, no stepping (derived-output CTRL)
      osc_torque(m, s, tg, grip, c, gn);
      conv = osc_converged(m, s, tg);
    }
    for (int t = 0; t < a.n_steps; t++) {
      PROF_T0
      csync(c, LK_TICK);
      PROF_ADD(PF_BARRIER)
      if (active) {
        osc_torque(m, s, tg, grip, c, gn);
      }
      PROF_ADD(PF_OSC)
      for (int k = 0; k < a.control_steps; k++) { step2<T, kPgs>(m, s, c, true, false, active); if (active) work += s.nefc * (2 + s.niter); step1(m, g, s, c, full, active); }
      if (active && osc_converged(m, s, tg)) conv = true;
    }
    if (active && c.tid == 0 && a.converged) a.converged[env] = conv ? 1 : 0;
  } else if (a.mode == MODE_SETTLE) {
    // PropPlacer settle loop (prop_initializer.py:245-257): arm + gripper held static, per-env early exit
    if (active) { for (int d = c.tid; d < m.ntd; d += c.nt) { s.hold_q[d] = s.qpos[d]; s.hold_v[d] = s.qvel[d]; } bsync(c); }
    int done = 0; bool fin = !active;
    for (int k = 0; k < a.n_steps; k++) {
      if (csync_and(c, fin)) break;
      step1(m, g, s, c, full, !fin); step2<T, kPgs>(m, s, c, true, true, !fin);
      if (!fin) {
        done = k + 1;
        T mx[2] = {0, 0};   // max |qvel|, |qacc| over ACTIVE prop dofs: computed redundantly by every lane (nv is tiny)
        for (int d = m.ntd; d < nv; d++) if (s.pactive[(d - m.ntd) / 6]) { const T v = fabs(s.qvel[d]), ac = fabs(s.qacc[d]); if (v > mx[0]) mx[0] = v; if (ac > mx[1]) mx[1] = ac; }
        bsync(c);
        if (mx[0] < a.settle_qvel_tol && mx[1] < a.settle_qacc_tol && done > a.settle_min_steps) fin = true;
      }
    }
    step1(m, g, s, c, full, active);
    if (active && c.tid == 0 && a.steps_done) a.steps_done[env] = done;
  }
  if (!active) return;
  bsync(c);
  if (!state_preloaded) { const T* head = s.qpos; for (int i = c.tid; i < REC_SIZE; i += c.nt) rec[i] = head[i]; }
  if (c.tid == 0 && a.status) a.status[env] |= s.flags;
  if (c.tid == 0 && a.work && (a.mode == MODE_PHYSICS || a.mode == MODE_OSC) && a.n_steps > 0) a.work[env] = work;
  if (a.out) { const int parts = a.mode == MODE_FORWARD ? (OUT_DYN | OUT_EFC) : ((a.mode == MODE_OSC && a.n_steps == 0) ? (a.out_mask | OUT_DYN | OUT_SITE) : a.out_mask); if (parts) write_out(m, g, s, c, a.out + (size_t)env * a.out_stride, parts); }
}

}  // namespace b2p
