#pragma once
#include "b2p_linalg.h"

namespace b2p {

// ------------------------------------------------------------------------------------------ mj_kinematics + comPos + comVel + rne (forward sweep)
template <typename T, bool C>
B2P_HDN void forward_sweep(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c) {
  if (c.tid == 0) {
    for (int k = 0; k < 3; k++) s.xpos[0][k] = 0;
    s.xquat[0][0] = 1; s.xquat[0][1] = 0; s.xquat[0][2] = 0; s.xquat[0][3] = 0;
    quat2mat(s.xmat[0], s.xquat[0]);
    for (int k = 0; k < 6; k++) { s.cvel[0][k] = 0; s.cacc[0][k] = 0; s.cfrc[0][k] = 0; }
    for (int k = 0; k < 3; k++) s.cacc[0][3 + k] = -m.gravity[k];
    for (int k = 0; k < 10; k++) s.crb[0][k] = 0;
  }
  for (int p = c.tid; p < m.nprop; p += c.nt) {     // free bodies: xpos = qpos[0:3], xquat = normalize(qpos[3:7])
    T* q = s.qpos + m.ntd + 7 * p;
    normquat(q + 3); cp3(s.ppos[p], q); quat2mat(s.pmat[p], q + 3);
  }
  bsync(c);
  // The tree is deep and narrow (the arm is a chain of 7, the gripper adds 3 more levels), so level-by-level passes are serial chains
  // run by one or two lanes.  A: the poses come from POINTER JUMPING instead -- every body first gets its transform relative to its
  // parent (one lane per body: nothing there depends on the parent's pose), then ceil(log2(depth)) rounds compose
  // T_b <- T_ptr(b) o T_b, ptr(b) <- ptr(ptr(b)) for all bodies at once: 4 rounds instead of 10 levels.  (Quaternions are normalised
  // once at the end instead of at every body: a difference of rounding order, ~1 ulp.)  C: the velocity / acceleration recursion
  // stays level by level (six-vectors, cheap).  B, D: inertia and force terms, one lane per body.
  for (int b = 1 + c.tid; b < m.ntb; b += c.nt) {
    const int d = b - 1;
    const T ha = T(0.5) * (s.qpos[d] - m.qpos0[d]), sn = sin(ha);
    const T qloc[4] = {cos(ha), sn * m.jnt_axis[d][0], sn * m.jnt_axis[d][1], sn * m.jnt_axis[d][2]};
    T q[4], rj[3];
    mulquat(q, m.tb_quat[b], qloc);                       // body frame relative to the parent frame
    rotquat(rj, q, m.jnt_pos[d]);
    for (int k = 0; k < 4; k++) s.xquat[b][k] = q[k];
    for (int k = 0; k < 3; k++) s.xpos[b][k] = m.tb_anchor[b][k] - rj[k];    // the joint anchor stays put: origin = anchor - R jnt_pos
    s.fkptr[b] = (unsigned char)m.tb_parent[b];
  }
  if (c.tid == 0) s.fkptr[0] = 0;
  bsync(c);
  for (int round = 0; round < m.fk_rounds; round++) {
    for (int b0 = 1; b0 < m.ntb; b0 += c.nt) {
      const int b = b0 + c.tid; const bool on = b < m.ntb;
      T q[4] = {1, 0, 0, 0}, t[3] = {0, 0, 0}; int pp = 0;
      if (on) {
        const int a = s.fkptr[b]; T rt[3];
        mulquat(q, s.xquat[a], s.xquat[b]);
        rotquat(rt, s.xquat[a], s.xpos[b]);
        for (int k = 0; k < 3; k++) t[k] = s.xpos[a][k] + rt[k];
        pp = s.fkptr[a];
      }
      bsync(c);                                           // every lane has read its ancestor's (transform, pointer) pair
      if (on) { for (int k = 0; k < 4; k++) s.xquat[b][k] = q[k]; for (int k = 0; k < 3; k++) s.xpos[b][k] = t[k]; s.fkptr[b] = (unsigned char)pp; }
      bsync(c);
    }
  }
  for (int b = 1 + c.tid; b < m.ntb; b += c.nt) { normquat(s.xquat[b]); quat2mat(s.xmat[b], s.xquat[b]); }
  bsync(c);
  for (int b = 1 + c.tid; b < m.ntb; b += c.nt) {       // joint anchor / axis in the world -> cdof (motion axis about the fixed reference point)
    const int p = m.tb_parent[b], d = b - 1;
    T xanchor[3], xaxis[3], off[3];
    mulmv(xanchor, s.xmat[p], m.tb_anchor[b]);
    for (int k = 0; k < 3; k++) xanchor[k] += s.xpos[p][k];
    mulmv(xaxis, s.xmat[p], m.tb_axis[b]);
    T* cd = s.cdof[d];
    cp3(cd, xaxis);
    for (int k = 0; k < 3; k++) off[k] = m.ref[k] - xanchor[k];
    cross3(cd + 3, xaxis, off);
  }
  bsync(c);
  for (int b = 1 + c.tid; b < m.ntb; b += c.nt) {   // B: inertia about the fixed reference point
    T xipos[3], qi[4], ximat[9], off[3];
    mulmv(xipos, s.xmat[b], m.tb_ipos[b]);
    for (int k = 0; k < 3; k++) off[k] = s.xpos[b][k] + xipos[k] - m.ref[k];
    mulquat(qi, s.xquat[b], m.tb_iquat[b]); quat2mat(ximat, qi);
    inert_com(s.crb[b], m.tb_inertia[b], ximat, off, m.tb_mass[b]);
  }
  for (int l = 0; l < m.nlevel; l++) {     // C: cvel = cvel_parent + cdof*qvel ; cacc = cacc_parent + (cvel_parent x cdof)*qvel
    for (int idx = m.level_start[l] + c.tid; idx < m.level_start[l + 1]; idx += c.nt) {
      const int b = m.level_body[idx], p = m.tb_parent[b], d = b - 1;
      const T* cd = s.cdof[d]; const T qv = s.qvel[d];
      T v[6], cdof_dot[6];
      for (int k = 0; k < 6; k++) v[k] = s.cvel[p][k];
      cross_motion(cdof_dot, v, cd);
      for (int k = 0; k < 6; k++) { s.cvel[b][k] = v[k] + cd[k] * qv; s.cacc[b][k] = s.cacc[p][k] + cdof_dot[k] * qv; }
    }
    bsync(c);
  }
  for (int b = 1 + c.tid; b < m.ntb; b += c.nt) {   // D: rne forward force  cfrc = I*cacc + cvel x* (I*cvel)
    T Ia[6], Iv[6], cf[6];
    mul_inert_vec(Ia, s.crb[b], s.cacc[b]); mul_inert_vec(Iv, s.crb[b], s.cvel[b]); cross_force(cf, s.cvel[b], Iv);
    for (int k = 0; k < 6; k++) s.cfrc[b][k] = Ia[k] + cf[k];
  }
  bsync(c);
  // geoms + sites
  for (int g = c.tid; g < m.ngeom; g += c.nt) {
    const int b = m.geom_body[g];
    const T *bp, *bm; T bq[4];
    if (b >= PROP_BODY0) { const int p = b - PROP_BODY0; bp = s.ppos[p]; bm = s.pmat[p]; for (int k = 0; k < 4; k++) bq[k] = s.qpos[m.ntd + 7 * p + 3 + k]; }
    else { bp = s.xpos[b]; bm = s.xmat[b]; for (int k = 0; k < 4; k++) bq[k] = s.xquat[b][k]; }
    T t[3], q[4];
    mulmv(t, bm, m.geom_pos[g]);
    for (int k = 0; k < 3; k++) s.gpos[g][k] = bp[k] + t[k];
    mulquat(q, bq, m.geom_quat[g]); quat2mat(s.gmat[g], q);
  }
  for (int i = c.tid; i < 2; i += c.nt) {
    const int b = m.site_body[i]; T t[3], q[4];
    mulmv(t, s.xmat[b], m.site_pos[i]);
    for (int k = 0; k < 3; k++) s.site_xpos[i][k] = s.xpos[b][k] + t[k];
    mulquat(q, s.xquat[b], m.site_quat[i]); quat2mat(s.site_xmat[i], q);
  }
  bsync(c);
}

// backward sweep: composite inertias (mj_crb) and rne force accumulation.  The sum over a body's subtree is taken directly -- one
// lane per body over the bodies whose ancestor mask holds it -- instead of parent-gathers-children level by level (nine dependent
// rounds with a barrier each on this chain-shaped tree).  (One lane per (body, component) was tried: more instructions in total --
// every item re-walks the mask -- and slower.)
template <typename T, bool C>
B2P_HD void backward_sweep(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c) {
  T acc[16];
  for (int b0 = 1; b0 < m.ntb; b0 += c.nt) {
    const int b = b0 + c.tid;
    if (b < m.ntb) {
      for (int k = 0; k < 16; k++) acc[k] = 0;
      for (int j = b; j < m.ntb; j++) if ((m.anc_mask[j] >> (b - 1)) & 1u) {       // descendants have larger ids (parent < child)
        for (int k = 0; k < 10; k++) acc[k] += s.crb[j][k];
        for (int k = 0; k < 6; k++) acc[10 + k] += s.cfrc[j][k];
      }
    }
    bsync(c);
    if (b < m.ntb) { for (int k = 0; k < 10; k++) s.crb[b][k] = acc[k]; for (int k = 0; k < 6; k++) s.cfrc[b][k] = acc[10 + k]; }
    bsync(c);
  }
}

// mass matrix (tree block dense, props diagonal), bias forces, passive forces
template <typename T, bool C>
B2P_HDN void dynamics_terms(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c) {
  const int n = m.ntd;
  // M_ij = cdof_j . (Ic_i cdof_i) for j on i's chain: the composite-inertia product once per dof (parked in cvel, dead since the force
  // pass), then one lane per element of the packed triangle
  for (int i = c.tid; i < n; i += c.nt) mul_inert_vec(s.cvel[i + 1], s.crb[i + 1], s.cdof[i]);
  bsync(c);
  for (int e = c.tid; e < (n * (n + 1)) / 2; e += c.nt) {
    const unsigned t = m.tri[e]; const int i = t >> 8, j = t & 255u;
    T v = 0;
    if (m.anc_mask[i + 1] & (1u << j)) {
      const T* buf = s.cvel[i + 1];
      for (int k = 0; k < 6; k++) v += s.cdof[j][k] * buf[k];
      if (i == j) v += m.dof_armature[i];
    }
    s.Mt[e] = v; s.Larm[e] = v;
  }
  for (int d = c.tid; d < m.nv; d += c.nt) {
    if (d < n) {
      T v = 0;
      for (int k = 0; k < 6; k++) v += s.cdof[d][k] * s.cfrc[d + 1][k];
      s.qfrc_bias[d] = v;
      s.qfrc_passive[d] = -m.dof_damping[d] * s.qvel[d] - m.jnt_stiffness[d] * (s.qpos[d] - m.qpos_spring[d]);
    } else {
      const int p = (d - n) / 6, k = (d - n) % 6; const T mass = s.pmass[p];
      if (k < 3) { s.Mdiag[d] = mass; s.qfrc_bias[d] = -mass * m.gravity[k]; }
      else {
        const T* w = s.qvel + n + 6 * p + 3; const T* I = s.pinertia[p];
        T Iw[3] = {I[0] * w[0], I[1] * w[1], I[2] * w[2]}, cr[3];
        cross3(cr, w, Iw);
        s.Mdiag[d] = I[k - 3]; s.qfrc_bias[d] = cr[k - 3];
      }
      s.qfrc_passive[d] = 0;
    }
  }
  bsync(c);
  sp_cholesky(s.Larm, m.sp_tree, m.tri, c);
}

// ------------------------------------------------------------------------------------------ actuation + smooth acceleration
template <typename T, bool C>
B2P_HDN void actuation_and_smooth(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c) {
  for (int d = c.tid; d < m.nv; d += c.nt) s.qfrc_act[d] = 0;
  bsync(c);
  for (int a = c.tid; a < m.nu; a += c.nt) {     // one lane per actuator: length, velocity, clamps, force
    T len = 0, vel = 0;
    for (int k = 0; k < m.act_trnnum[a]; k++) { const int d = m.act_trndof[a][k]; len += m.act_trncoef[a][k] * s.qpos[d]; vel += m.act_trncoef[a][k] * s.qvel[d]; }
    T u = s.ctrl[a];
    if (m.act_ctrllimited[a]) u = clampT(u, m.act_ctrlrange[a][0], m.act_ctrlrange[a][1]);
    T f = m.act_gain[a] * u + m.act_bias[a][0] + m.act_bias[a][1] * len + m.act_bias[a][2] * vel;
    if (m.act_forcelimited[a]) f = clampT(f, m.act_forcerange[a][0], m.act_forcerange[a][1]);
    s.act_force[a] = f;
  }
  bsync(c);
  if (c.tid == 0)                                 // moment' * force, in actuator order (two actuators may share a dof)
    for (int a = 0; a < m.nu; a++) for (int k = 0; k < m.act_trnnum[a]; k++) s.qfrc_act[m.act_trndof[a][k]] += m.act_trncoef[a][k] * s.act_force[a];
  bsync(c);
  for (int d = c.tid; d < m.nv; d += c.nt) {
    bool active = d < m.ntd || s.pactive[(d - m.ntd) / 6];
    const T f = active ? s.qfrc_passive[d] - s.qfrc_bias[d] + s.qfrc_act[d] : T(0);
    s.qfrc_smooth[d] = f;
    s.qacc_smooth[d] = d < m.ntd ? f : f / s.Mdiag[d];
  }
  bsync(c);
  sp_solve(s.Larm, m.sp_tree, s.qacc_smooth, 0, m.ntd, c);
}

// ------------------------------------------------------------------------------------------ integrator (mj_implicit / mj_Euler + mj_advance)
template <typename T, bool C>
B2P_HDN void integrate(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c, bool freeze_tree) {
  const int n = m.ntd; const T h = m.timestep;
  // M - h * d(qfrc_passive + qfrc_actuator)/d(qvel): the velocity-derivative part is a model constant (joint damping
  // on the diagonal; for implicitfast the affine actuator's -bias_vel * moment' moment inside M's sparsity), built on the host
  for (int e = c.tid; e < (n * (n + 1)) / 2; e += c.nt) { const unsigned t = m.tri[e]; s.H[e] = s.Mt[e] + m.impl_hD[e]; }
  for (int d = c.tid; d < m.nv; d += c.nt) {
    bool active = d < n || s.pactive[(d - n) / 6];
    const T f = active ? s.qfrc_smooth[d] + s.qfrc_con[d] : T(0);
    s.tmpv[d] = d < n ? f : f / s.Mdiag[d];
  }
  bsync(c);
  sp_cholesky(s.H, m.sp_tree, m.tri, c);
  sp_solve(s.H, m.sp_tree, s.tmpv, 0, n, c);
  for (int d = c.tid; d < m.nv; d += c.nt) { s.qvel[d] += h * s.tmpv[d]; s.qacc_ws[d] = s.qacc[d]; }
  bsync(c);
  for (int d = c.tid; d < n; d += c.nt) s.qpos[d] += h * s.qvel[d];
  for (int p = c.tid; p < m.nprop; p += c.nt) {
    T* q = s.qpos + n + 7 * p; const T* v = s.qvel + n + 6 * p;
    for (int k = 0; k < 3; k++) q[k] += h * v[k];
    T w[3] = {v[3], v[4], v[5]}; const T ang = h * normalize3(w), sn = sin(T(0.5) * ang);
    T qr[4] = {cos(T(0.5) * ang), sn * w[0], sn * w[1], sn * w[2]};
    mulquat(q + 3, q + 3, qr); normquat(q + 3);
  }
  if (c.tid == 0) s.time += h;
  bsync(c);
  if (freeze_tree) {   // dm_control JointStaticIsolator: non-prop joints restored after the step (App. B item 18)
    for (int d = c.tid; d < n; d += c.nt) { s.qpos[d] = s.hold_q[d]; s.qvel[d] = s.hold_v[d]; }
    bsync(c);
  }
}


}  // namespace b2p
