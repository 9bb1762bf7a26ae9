"""Generate tests/golden/osc_golden.npz: known-answer vectors for the OSC half of the hot path.

The reference's OSC (mujoco_controllers/osc.py:99-238) is numpy on top of five MuJoCo queries (site pose, mj_jacSite,
mj_fullM, qfrc_bias, actuator_moment).  MuJoCo is not installable here, so the five query results are taken from the
CPU oracle at seeded random states -- the site Jacobian is additionally cross-checked against central finite
differences of the site pose, which does not depend on any Jacobian code -- and THE REFERENCE'S FORMULAE are then
evaluated with the same numpy.linalg calls the reference makes (inv, det, pinv(rcond=1e-2), pinv of the dense
actuator-moment matrix).  The stored torque is therefore "reference osc.py arithmetic on given MuJoCo-style inputs";
the oracle's and the kernels' own OSC code (Cholesky / Jacobi based, no numpy) must reproduce it.

  python tests/golden/make_osc_golden.py        # rewrites osc_golden.npz next to this file
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import scenarios as sc  # noqa: E402
from orc import Oracle  # noqa: E402

GAINS = {"position": {"kp": 350.0, "kd": 20.0}, "orientation": {"kp": 500.0, "kd": 100.0}, "nullspace": {"kp": 200.0, "kd": 30.0}}
NULLSPACE = np.array([0, -0.785, 0, -2.356, 0, 1.571, 0.785])
ARM = np.arange(7)
SITE = 3   # attachment_site (names.json "site" order)


def mat2quat(m):          # mju_mat2Quat
    m = m.reshape(3, 3)
    if m[0, 0] + m[1, 1] + m[2, 2] > 0:
        w = 0.5 * np.sqrt(1 + m[0, 0] + m[1, 1] + m[2, 2]); q = [w, 0.25 * (m[2, 1] - m[1, 2]) / w, 0.25 * (m[0, 2] - m[2, 0]) / w, 0.25 * (m[1, 0] - m[0, 1]) / w]
    elif m[0, 0] > m[1, 1] and m[0, 0] > m[2, 2]:
        x = 0.5 * np.sqrt(1 + m[0, 0] - m[1, 1] - m[2, 2]); q = [0.25 * (m[2, 1] - m[1, 2]) / x, x, 0.25 * (m[0, 1] + m[1, 0]) / x, 0.25 * (m[0, 2] + m[2, 0]) / x]
    elif m[1, 1] > m[2, 2]:
        y = 0.5 * np.sqrt(1 - m[0, 0] + m[1, 1] - m[2, 2]); q = [0.25 * (m[0, 2] - m[2, 0]) / y, 0.25 * (m[0, 1] + m[1, 0]) / y, y, 0.25 * (m[1, 2] + m[2, 1]) / y]
    else:
        z = 0.5 * np.sqrt(1 - m[0, 0] - m[1, 1] + m[2, 2]); q = [0.25 * (m[1, 0] - m[0, 1]) / z, 0.25 * (m[0, 2] + m[2, 0]) / z, 0.25 * (m[1, 2] + m[2, 1]) / z, z]
    q = np.array(q); return q / np.linalg.norm(q)


def mulquat(a, b):        # mju_mulQuat
    return np.array([a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3], a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
                     a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1], a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]])


def reference_osc(xpos, xmat, J, M, qpos, qvel, bias, moment, tg):
    """osc.py:99-115, 132-150, 160-238 -- same numpy calls, same order."""
    M_inv = np.linalg.inv(M)
    Mx_inv = np.dot(J, np.dot(M_inv, J.T))
    branch = int(abs(np.linalg.det(Mx_inv)) >= 1e-2)
    Mx = np.linalg.inv(Mx_inv) if branch else np.linalg.pinv(Mx_inv, rcond=1e-2)
    quat = mat2quat(xmat); quat /= np.linalg.norm(quat)
    qd = tg[6:10] / np.linalg.norm(tg[6:10])                 # set_target normalises (osc.py:95)
    def ori_err(q, q_des):
        qc = np.array([q[0], -q[1], -q[2], -q[3]]); qc /= np.linalg.norm(qc)
        e = mulquat(q_des, qc)
        return e[1:] * np.sign(e[0])
    dx, dw = J[:3] @ qvel, J[3:] @ qvel
    position_pd = GAINS["position"]["kp"] * np.clip(tg[0:3] - xpos, -0.05, 0.05) + GAINS["position"]["kd"] * (tg[3:6] - dx)
    orientation_pd = GAINS["orientation"]["kp"] * ori_err(quat, qd) + GAINS["orientation"]["kd"] * (tg[10:13] - dw)
    tau = J.T @ Mx @ np.hstack([position_pd, orientation_pd])
    null_err = GAINS["nullspace"]["kp"] * (NULLSPACE - qpos) + GAINS["nullspace"]["kd"] * (np.zeros(7) - qvel)
    null_jac = np.linalg.inv(M) @ J.T @ Mx
    tau = tau + (np.eye(7) - J.T @ null_jac.T) @ null_err
    tau = tau + bias
    inv = np.linalg.pinv(moment)[ARM, :][:, ARM]
    tau = tau @ inv
    err = ori_err(quat, qd)
    conv = (np.linalg.norm(xpos - tg[0:3]) < 5e-3) and (np.max(np.abs(err)) < 68e-3)
    return tau, branch, bool(conv)


def site_jacobian(o):
    """mj_jacSite from cdof (spatial axes about the subtree-root com, [ang; lin]):  jacr = ang, jacp = lin + ang x (p - com)."""
    cdof, com, p = o.cdof[:7], o.subtree_com[1], o.site_xpos[SITE]
    jacr = cdof[:, 0:3].T.copy()
    jacp = (cdof[:, 3:6] + np.cross(cdof[:, 0:3], p - com)).T
    return np.vstack([jacp, jacr])


def fd_site_jacobian(o, q0, eps=1e-6):
    J = np.zeros((6, 7))
    for i in range(7):
        pp, mm = [], []
        for sgn, acc in ((1, pp), (-1, mm)):
            o.qpos[:7] = q0; o.qpos[i] += sgn * eps; o.forward()
            acc.append(o.site_xpos[SITE].copy()); acc.append(o.site_xmat[SITE].reshape(3, 3).copy())
        J[:3, i] = (pp[0] - mm[0]) / (2 * eps)
        dR = (pp[1] - mm[1]) / (2 * eps) @ pp[1].T          # dR/dq R^T = [w]x
        J[3:, i] = [dR[2, 1], dR[0, 2], dR[1, 0]]
    o.qpos[:7] = q0; o.forward()
    return J


def moment_matrix():
    """actuator_moment of the compiled scene, dense [nu, nv]: unit rows on dofs 0..6, 0.5 on the two driver dofs."""
    A = np.zeros((sc.NU, sc.NV))
    A[np.arange(7), np.arange(7)] = 1.0
    A[7, 7] = 0.5; A[7, 11] = 0.5
    return A


def main():
    rng = np.random.default_rng(20261104)
    o = Oracle(sc.default_blob())
    lo = np.array([-2.8, -1.7, -2.8, -3.0, -2.8, 0.0, -2.8]); hi = np.array([2.8, 1.7, 2.8, -0.1, 2.8, 3.7, 2.8])
    n = 48
    rec = {k: [] for k in ("qpos", "qvel", "target", "xpos", "xmat", "J", "M", "bias", "tau", "branch", "converged")}
    st, size, active = sc.make_states(1, seed=1)
    worst_fd = 0.0
    for i in range(n):
        sc.oracle_load(o, st[0], size[0], active[0])
        if i % 4 == 0:     # near the stretched-out singularity: exercises the pinv branch
            q = np.array([rng.uniform(-1, 1), rng.uniform(-0.2, 0.2), 0, rng.uniform(-0.15, -0.08), 0, rng.uniform(0, 0.2), 0])
        elif i % 4 == 1:   # around home (the states the rollouts visit)
            q = sc.ARM_HOME + 0.3 * (rng.random(7) - 0.5)
        else:
            q = lo + (hi - lo) * rng.random(7)
        o.qpos[:7] = q; o.qvel[:7] = rng.normal(0, 0.5, 7); o.forward()
        J = site_jacobian(o)
        worst_fd = max(worst_fd, np.abs(J - fd_site_jacobian(o, q.copy())).max())
        o.forward()
        M = o.qm[:7, :7].copy(); bias = o.qfrc_bias[:7].copy()
        xpos, xmat = o.site_xpos[SITE].copy(), o.site_xmat[SITE].copy()
        tg = np.zeros(13)
        if i % 6 == 5:     # target at the current pose -> converged, tau = nullspace + bias terms only
            tg[0:3] = xpos; tg[6:10] = mat2quat(xmat)
        else:
            tg[0:3] = xpos + rng.normal(0, 0.08, 3); tg[3:6] = rng.normal(0, 0.05, 3)
            tg[6:10] = mulquat(mat2quat(xmat), np.r_[1.0, rng.normal(0, 0.2, 3)]) * 1.7; tg[10:13] = rng.normal(0, 0.1, 3)
        tau, branch, conv = reference_osc(xpos, xmat, J, M, o.qpos[:7].copy(), o.qvel[:7].copy(), bias, moment_matrix(), tg)
        for k, v in (("qpos", o.qpos[:7].copy()), ("qvel", o.qvel[:7].copy()), ("target", tg), ("xpos", xpos), ("xmat", xmat), ("J", J), ("M", M),
                     ("bias", bias), ("tau", tau), ("branch", branch), ("converged", conv)):
            rec[k].append(v)
    assert worst_fd < 1e-6, worst_fd
    out = {k: np.array(v) for k, v in rec.items()}
    out["scene_state"] = st[0]; out["scene_size"] = size[0]; out["scene_active"] = active[0]
    print("finite-difference check of the site Jacobian: max |J - J_fd| = %.2e" % worst_fd)
    print("branches: inverse %d, pinv %d; converged %d" % (out["branch"].sum(), n - out["branch"].sum(), out["converged"].sum()))
    np.savez_compressed(os.path.join(HERE, "osc_golden.npz"), **out)


if __name__ == "__main__":
    main()
