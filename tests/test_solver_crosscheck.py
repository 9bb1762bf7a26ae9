"""Newton (primal) against PGS (dual) on identical constraint sets (SURVEY 8c fixture 3; north star: "PGS").

The two solvers share the rows (J, aref, R) but nothing else: Newton minimises 1/2 (a - a0)'M(a - a0) + s(Ja - aref) with the
R-scaled elliptic cone and contact.mu; PGS minimises the dual 1/2 f'(J M^-1 J' + R)f + f'(J a0 - aref) over the TRUE friction
cones.  They agree only if R, D = 1/R, the friction-row scaling by impratio, contact.mu and the three cone zones of the primal
cost are mutually consistent -- a check of the objective that does not need MuJoCo.  Tolerance: 1e-6 relative on qacc (fp64).
"""
import numpy as np
import pytest

import scenarios as sc
from orc import Oracle


def _both(o, st, size, active, ctrl):
    out = {}
    for solver, its, tol in ((0, 100, 1e-12), (1, 50000, 1e-17)):
        sc.oracle_load(o, st, size, active); o.set_option("full_contacts", 0); o.ctrl[:sc.NU] = ctrl
        o.set_option("solver", solver); o.set_option("iterations", its); o.set_option("tolerance", tol)
        o.forward()
        out[solver] = (o.qacc[:sc.NV].copy(), o.efc_force.copy(), o.dim("solver_iter"), o.dim("nefc"), o.contacts())
    o.set_option("solver", 0); o.set_option("iterations", 100); o.set_option("tolerance", 1e-8)
    return out


def _scenes():
    """(label, state, size, active, ctrl): resting, landing, sliding, grasping, stacked, and arm-flailing states, all produced by
    stepping the oracle with its default (Newton) solver"""
    blob = sc.default_blob()
    o = Oracle(blob)
    out = []
    st, size, active = sc.make_states(2, seed=7)
    out.append(("rest", st[0], size[0], active[0], np.zeros(sc.NU)))
    st, size, active = sc.make_states(1, seed=11, settled=False)
    sc.oracle_load(o, st[0], size[0], active[0]); o.set_option("full_contacts", 0)
    for _ in range(75):
        o.step()
    out.append(("landing", sc.oracle_state(o), size[0], active[0], np.zeros(sc.NU)))
    for _ in range(200):
        o.step()
    o.qvel[15] = 0.4; o.qvel[15 + 6 + 1] = -0.3
    out.append(("sliding", sc.oracle_state(o), size[0], active[0], np.zeros(sc.NU)))
    for kind in ("grasp", "stack", "grasp_stack"):
        s, z, a = sc.coupled_states(kind)
        sc.oracle_load(o, s[0], z[0], a[0]); o.set_option("full_contacts", 0); o.step1()
        c = np.zeros(sc.NU); c[:7] = o.qfrc_bias[:7]; c[7] = 255.0
        o.ctrl[:sc.NU] = c
        for _ in range(40):
            o.step()
        out.append((kind, sc.oracle_state(o), z[0], a[0], c))
    st, size, active = sc.make_states(1, seed=1234)
    sc.oracle_load(o, st[0], size[0], active[0]); o.set_option("full_contacts", 0); o.step1(); bias = o.qfrc_bias[:7].copy()
    for k in range(400):
        o.ctrl[:sc.NU] = sc.halton_ctrl(1, k // 100, bias)[0]
        o.step()
    out.append(("halton", sc.oracle_state(o), size[0], active[0], sc.halton_ctrl(1, 3, bias)[0]))
    return out


@pytest.fixture(scope="module")
def scenes():
    return _scenes()


def test_newton_and_pgs_agree(scenes):
    o = Oracle(sc.default_blob())
    seen_cone = False
    for label, st, size, active, ctrl in scenes:
        r = _both(o, st, size, active, ctrl)
        qn, fn, itn, nefc, con = r[0]; qp, fp, itp, _, _ = r[1]
        assert itp < 50000, label                                   # PGS met its own tolerance
        rel = np.abs(qn - qp).max() / max(1.0, np.abs(qn).max())
        assert rel < 1e-6, (label, rel, nefc, itn, itp)
        assert np.abs(fn - fp).max() < 1e-5 * max(1.0, np.abs(fn).max()), label
        # dual feasibility of both force vectors in the TRUE cone: f_n >= 0, |f_t| <= mu f_n (mu = 1 for prop pairs, 0.7 / 0.6 pad pairs)
        ncr = 3 * int((con["exclude"] == 0).sum())
        for f in (fn, fp):
            fc = f[nefc - ncr:nefc].reshape(-1, 3)
            assert (fc[:, 0] >= -1e-9).all(), label
            assert (np.hypot(fc[:, 1], fc[:, 2]) <= 1.0 * fc[:, 0] + 1e-6).all(), label
            seen_cone |= bool(((np.hypot(fc[:, 1], fc[:, 2]) > 0.999 * fc[:, 0]) & (fc[:, 0] > 1e-3)).any())
    assert seen_cone                                                # at least one contact slides (force on the cone boundary)


def test_pgs_steps_like_newton(scenes):
    """40 steps with each solver from the grasp state: trajectories stay within 1e-7"""
    label, st, size, active, ctrl = next(s for s in scenes if s[0] == "grasp")
    o = Oracle(sc.default_blob())
    res = []
    for solver, its, tol in ((0, 100, 1e-12), (1, 50000, 1e-17)):
        sc.oracle_load(o, st, size, active); o.set_option("full_contacts", 0); o.ctrl[:sc.NU] = ctrl
        o.set_option("solver", solver); o.set_option("iterations", its); o.set_option("tolerance", tol)
        for _ in range(40):
            o.step()
        res.append(o.qpos[:sc.NQ].copy())
    assert np.abs(res[0] - res[1]).max() < 1e-7


def test_zero_duality_gap_at_the_newton_solution(scenes):
    """numpy restatement of BOTH objectives from the exported rows: the primal cost (Gauss term + the three cone zones) at Newton's
    qacc plus the dual cost 1/2 f'(J M^-1 J' + R) f + f'(J a0 - aref) at Newton's forces is zero (strong duality), the forces lie in
    the true friction cones, and qacc = a0 + M^-1 J' f."""
    from mujoco_robot_environments_b200.compile import blob as B
    _, _, arr = B.unpack(sc.default_blob())

    def fric(g1, g2):
        p1, p2 = arr["geom_priority"][g1], arr["geom_priority"][g2]
        if p1 != p2:
            return arr["geom_friction"][g1 if p1 > p2 else g2][0]
        return max(arr["geom_friction"][g1][0], arr["geom_friction"][g2][0])
    o = Oracle(sc.default_blob())
    for label, st, size, active, ctrl in scenes:
        sc.oracle_load(o, st, size, active); o.set_option("full_contacts", 0); o.ctrl[:sc.NU] = ctrl
        o.set_option("tolerance", 1e-12); o.forward(); o.set_option("tolerance", 1e-8)
        nefc = o.dim("nefc")
        J, R, aref, M, a0, qn, fn = o.efc_j.copy(), o.efc_r.copy(), o.efc_aref.copy(), o.qm.copy(), o.qacc_smooth.copy(), o.qacc.copy(), o.efc_force.copy()
        D = 1 / R
        con = o.contacts()
        pairs = con["geom"][con["exclude"] == 0]
        r0 = nefc - 3 * len(pairs)
        act = np.concatenate([np.ones(15, bool), np.repeat(np.asarray(active, bool), 6)])
        Mi = np.zeros_like(M); Mi[np.ix_(act, act)] = np.linalg.inv(M[np.ix_(act, act)])
        jar = J @ qn - aref
        da = (qn - a0)[act]
        primal = 0.5 * da @ M[np.ix_(act, act)] @ da
        for i in range(r0):
            if i < 7 or jar[i] < 0:                     # 7 equality rows, then active limit rows
                primal += 0.5 * D[i] * jar[i] ** 2
        for k, (g1, g2) in enumerate(pairs):
            i, fr = r0 + 3 * k, fric(g1, g2)
            mu = fr * np.sqrt(R[i + 1] / R[i])
            N, T = jar[i] * mu, np.hypot(jar[i + 1] * fr, jar[i + 2] * fr)
            if N >= mu * T:
                pass
            elif mu * N + T <= 0:
                primal += 0.5 * (D[i:i + 3] * jar[i:i + 3] ** 2).sum()
            else:
                primal += 0.5 * D[i] / (mu * mu * (1 + mu * mu)) * (N - mu * T) ** 2
            assert fn[i] >= -1e-9 and np.hypot(fn[i + 1], fn[i + 2]) <= fr * fn[i] + 1e-7, (label, k)
        A = J @ Mi @ J.T + np.diag(R)
        dual = 0.5 * fn @ A @ fn + fn @ (J @ a0 - aref)
        assert abs(primal + dual) < 1e-7 * max(1.0, abs(primal)), (label, primal, dual)
        assert np.abs((a0 + Mi @ J.T @ fn - qn)[act]).max() < 1e-6 * max(1.0, np.abs(qn).max()), label


def _pgs_blob():
    import blobtools as bt
    return bt.surgery(sc.default_blob(), opts=dict(iterations=50000.0, tolerance=1e-17))


def test_kernel_pgs_agrees_with_newton_on_the_host_build(scenes):
    """The kernel's own PGS (csrc/b2p_pgs.h, matrix-free: b2p_set_solver(B2P_SOLVER_PGS)) compiled for the host, run to its tolerance on
    every scene, against the ORACLE's Newton: qacc within 1e-6 relative.  The product default stays Newton; the GPU leg is
    tests/test_gpu_parity.py::test_pgs_solver_on_gpu."""
    from hostemu import MODE_FORWARD, HostEmu, out_slices
    S = out_slices()
    o = Oracle(sc.default_blob()); o.set_option("tolerance", 1e-12)
    e = HostEmu(_pgs_blob(), np.float64)
    for label, st, size, active, ctrl in scenes:
        sc.oracle_load(o, st, size, active); o.set_option("full_contacts", 0); o.ctrl[:sc.NU] = ctrl; o.forward()
        r = e.run(MODE_FORWARD, st[None].copy(), ctrl=ctrl[None], prop_size=size[None], prop_active=active[None], solver=1)
        out = r["out"][0]
        qn, qp = o.qacc[:sc.NV], out[S["qacc"]][:sc.NV]
        assert int(out[S["niter"]][0]) < 50000 and (r["status"][0] & 64) == 0, label
        assert int(out[S["nefc"]][0]) == o.dim("nefc"), label
        rel = np.abs(qn - qp).max() / max(1.0, np.abs(qn).max())
        assert rel < 1e-6, (label, rel)
        fo = out[S["efc_force"]][:o.dim("nefc")]
        assert np.abs(fo - o.efc_force).max() < 1e-5 * max(1.0, np.abs(o.efc_force).max()), label
