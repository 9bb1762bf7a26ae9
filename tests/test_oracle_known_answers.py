"""Analytic known-answer tests that pin the CPU oracle (oracle/b2p_oracle.c).

The reference holds no golden vectors for this path and MuJoCo is not installable here (SURVEY.md 8c:
"parity unpinned"), so the oracle is anchored on physics whose answer is known in closed form or can be
re-derived independently in numpy (finite-difference Jacobians, potential-energy gradients, KKT residuals).
"""
import numpy as np
import pytest

import scenarios as sc
from orc import Oracle
from mujoco_robot_environments_b200.compile import blob as B

G = 9.8
H = 1e-3


@pytest.fixture(scope="module")
def model():
    blob = sc.default_blob()
    counts, opts, arrays = B.unpack(blob)
    return blob, counts, opts, arrays


def _fresh(blob, seed=1, settled=True, jitter=0.15):
    o = Oracle(blob)
    st, size, active = sc.make_states(1, seed=seed, settled=settled, jitter=jitter)
    sc.oracle_load(o, st[0], size[0], active[0])
    return o, st[0], size[0], active[0]


def _q2m(q):
    w, x, y, z = q
    return np.array([[w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z]])


def test_free_fall_is_exact_semi_implicit_euler(model):
    """A prop far above everything (outside every margin): v_n = -g h n, z_n = z0 - g h^2 n(n+1)/2 (mj_Euler/implicitfast
    are identical for an undamped free body; App. B item 15)."""
    blob = model[0]
    o, st, size, active = _fresh(blob)
    q = 15
    o.qpos[q:q + 3] = (0.5, 0.0, 3.0)
    o.qpos[q + 3:q + 7] = (1, 0, 0, 0)
    n = 200
    for _ in range(n):
        o.step()
    assert abs(o.qvel[15 + 2] - (-G * H * n)) < 1e-12
    assert abs(o.qpos[q + 2] - (3.0 - G * H * H * n * (n + 1) / 2)) < 1e-12
    assert np.abs(o.qpos[q:q + 2] - (0.5, 0.0)).max() < 1e-14


def test_free_spin_keeps_unit_quaternion_and_angular_momentum(model):
    """Torque-free cube (equal principal inertias): body angular velocity stays constant, |quat| = 1."""
    blob = model[0]
    o, *_ = _fresh(blob)
    q = 15
    o.qpos[q:q + 3] = (0.5, 0.0, 3.0)
    o.qpos[q + 3:q + 7] = (1, 0, 0, 0)
    o.qvel[15 + 3:15 + 6] = (1.0, -2.0, 0.5)
    for _ in range(100):
        o.step()
    assert np.abs(o.qvel[18:21] - (1.0, -2.0, 0.5)).max() < 1e-12
    assert abs(np.linalg.norm(o.qpos[q + 3:q + 7]) - 1) < 1e-14
    # rotation angle = |w| * t
    ang = 2 * np.arccos(np.clip(o.qpos[q + 3], -1, 1))
    assert abs(ang - np.linalg.norm([1.0, -2.0, 0.5]) * 0.1) < 1e-9


def test_resting_cube_normal_force_balances_weight(model):
    """Settled cube on the table: contact normal forces sum to m g; qacc ~ 0; penetration tiny and positive."""
    blob, counts, opts, arrays = model
    o, st, size, active = _fresh(blob, seed=2)
    o.set_option("full_contacts", 0)
    for _ in range(600):
        o.step_frozen_arm()
    o.forward()
    nprop = int(active.sum())
    assert np.abs(o.qacc[15:15 + 6 * nprop]).max() < 1e-3
    assert np.abs(o.qvel[15:15 + 6 * nprop]).max() < 1e-4
    # qfrc_constraint on each active prop's z dof == +m g
    for p in range(nprop):
        assert abs(o.qfrc_constraint[15 + 6 * p + 2] - 0.1 * G) < 1e-4
        z = o.qpos[15 + 7 * p + 2]
        pen = 0.4 + size[p][2] - z
        assert 0 < pen < 5e-4


def test_sliding_cube_friction_sits_on_cone_boundary(model):
    """mu = 1 (prop friction wins by priority, props.py:194-210): the step after a horizontal kick the cube slides, so the
    net contact force sits on the friction-cone boundary |f_t| = mu f_n, opposing the motion.  (A 3 cm cube with mu = 1
    is at its tipping threshold, so a longer slide tumbles; only the first step has a closed-form answer.)"""
    blob = model[0]
    o, st, size, active = _fresh(blob, seed=2)
    o.set_option("full_contacts", 0)
    for _ in range(400):
        o.step_frozen_arm()
    o.qvel[15] = 0.5
    o.step_frozen_arm()
    fx, fy, fz = o.qfrc_constraint[15:18]
    assert fx < 0 and fz > 0
    assert 0.98 < np.hypot(fx, fy) / fz <= 1.0 + 1e-6
    # a slow nudge is arrested by static friction: no creep
    o2, *_ = _fresh(blob, seed=2)
    o2.set_option("full_contacts", 0)
    for _ in range(400):
        o2.step_frozen_arm()
    o2.qvel[15] = 0.005
    for _ in range(30):
        o2.step_frozen_arm()
    assert abs(o2.qvel[15]) < 0.1 * 0.005     # soft friction rows: exponential decay, not a hard stop


def test_gravity_bias_is_potential_energy_gradient(model):
    """qfrc_bias(q, v = 0) = dPE/dq with PE = sum_b m_b g z_com,b — pins FK + inertial frames + RNE gravity jointly."""
    blob, counts, opts, arrays = model
    o, st, *_ = _fresh(blob, seed=3, jitter=0.6)
    o.qpos[7:15] = [0.3, -0.2, 0.25, 0.1, 0.3, -0.2, 0.25, 0.1]
    mass = arrays["body_mass"]

    def pe():
        o.forward()
        return float((mass * G * o.xipos[:, 2]).sum())
    o.forward()
    bias = o.qfrc_bias[:15].copy()
    eps = 1e-6
    for d in range(15):
        q0 = o.qpos[d]
        o.qpos[d] = q0 + eps; ep = pe()
        o.qpos[d] = q0 - eps; em = pe()
        o.qpos[d] = q0
        assert abs((ep - em) / (2 * eps) - bias[d]) < 2e-6 * max(1.0, abs(bias[d]))


def test_mass_matrix_from_independent_jacobians(model):
    """qM (CRBA) = sum_b m Jc'Jc + Jw' R I R' Jw + armature, with Jc / Jw from finite differences of xipos / xmat."""
    blob, counts, opts, arrays = model
    o, st, *_ = _fresh(blob, seed=4, jitter=0.6)
    o.qpos[7:15] = [0.3, -0.2, 0.25, 0.1, 0.3, -0.2, 0.25, 0.1]
    o.forward()
    M = o.qm[:15, :15].copy()
    nb = counts["nbody"]
    mass, inertia, iquat = arrays["body_mass"], arrays["body_inertia"], arrays["body_iquat"]
    xmat0 = o.xmat.reshape(nb, 3, 3).copy()
    eps = 1e-6
    Jc, Jw = np.zeros((nb, 3, 15)), np.zeros((nb, 3, 15))
    for d in range(15):
        q0 = o.qpos[d]
        o.qpos[d] = q0 + eps; o.forward(); xp, Rp = o.xipos.copy(), o.xmat.reshape(nb, 3, 3).copy()
        o.qpos[d] = q0 - eps; o.forward(); xm, Rm = o.xipos.copy(), o.xmat.reshape(nb, 3, 3).copy()
        o.qpos[d] = q0
        Jc[:, :, d] = (xp - xm) / (2 * eps)
        for b in range(nb):
            W = (Rp[b] - Rm[b]) / (2 * eps) @ xmat0[b].T
            Jw[b, :, d] = (W[2, 1], W[0, 2], W[1, 0])
    o.forward()
    Mref = np.diag(arrays["dof_armature"][:15]).astype(float)
    for b in range(1, nb):
        if arrays["body_dofadr"][b] >= 15:
            continue
        Ri = xmat0[b] @ _q2m(iquat[b])
        Iw = Ri @ np.diag(inertia[b]) @ Ri.T
        Mref += mass[b] * Jc[b].T @ Jc[b] + Jw[b].T @ Iw @ Jw[b]
    assert np.abs(M - Mref).max() < 1e-6
    assert np.abs(M - M.T).max() < 1e-14
    assert np.linalg.eigvalsh(M).min() > 0


def test_smooth_acceleration_solves_M_qacc_equals_force(model):
    blob = model[0]
    o, *_ = _fresh(blob, seed=5)
    o.qvel[:15] = np.linspace(-0.5, 0.5, 15)
    o.ctrl[:] = [5, -3, 2, 1, 0.5, -0.5, 0.2, 120]
    o.forward()
    o2, st, size, active = _fresh(blob, seed=5)
    act = np.concatenate([np.ones(15, bool), np.repeat(active.astype(bool), 6)])     # parked (inactive) props are frozen
    r = (o.qm @ o.qacc_smooth - o.qfrc_smooth)[act]
    assert np.abs(r).max() < 1e-9 * max(1, np.abs(o.qfrc_smooth).max())
    assert np.abs(o.qfrc_smooth - (o.qfrc_passive - o.qfrc_bias + o.qfrc_actuator))[act].max() < 1e-12


def test_solver_kkt_residual_and_cone_feasibility(model):
    """At the Newton solution: M (qacc - qacc_smooth) = J' f ; every contact force inside its (scaled) friction cone,
    normal forces >= 0; limit forces >= 0."""
    blob = model[0]
    o, st, size, active = _fresh(blob, seed=6, settled=False)
    o.set_option("full_contacts", 0)
    for _ in range(60):    # cubes fall 2 cm and hit the table
        o.step()
    o.qvel[15] = 0.2
    o.forward()
    nefc = o.dim("nefc")
    assert nefc > 7
    J, f = o.efc_j, o.efc_force
    r = o.qm @ (o.qacc - o.qacc_smooth) - J.T @ f
    assert np.abs(r).max() < 1e-6
    assert np.abs(o.qfrc_constraint - J.T @ f).max() < 1e-10
    # rows: 7 equality, then limits, then contacts in triples
    ncon_rows = 3 * sum(1 for d, e in zip(o.contacts()["dist"], o.contacts()["exclude"]) if e == 0)
    fc = f[nefc - ncon_rows:].reshape(-1, 3)
    assert (fc[:, 0] >= -1e-12).all()
    mu = 1.0
    assert (np.hypot(fc[:, 1], fc[:, 2]) <= mu * fc[:, 0] + 1e-9).all()


def test_gripper_closure_and_actuator(model):
    """Closing the 2F-85 for 0.5 s: the two connect equalities and the driver-coupling equality stay closed
    (residual < 1e-3), and the finger actuator force is clamped to forcerange +-1.5 (robotiq_2f85.py:41-48)."""
    blob = model[0]
    o, *_ = _fresh(blob, seed=7)
    o.set_option("full_contacts", 0)
    o.forward(); o.ctrl[:7] = o.qfrc_bias[:7]; o.ctrl[7] = 255.0
    worst = 0.0
    for _ in range(500):
        o.step()
        worst = max(worst, float(np.abs(o.efc_pos[:7]).max()))
        assert abs(o.act_force[7]) <= 1.5 + 1e-12
    assert worst < 1e-3
    assert o.qpos[7] > 0.3 and abs(o.qpos[7] - o.qpos[11]) < 1e-3    # drivers closed and mirrored


def test_joint_limit_row_activates_only_past_range(model):
    blob, counts, opts, arrays = model
    o, *_ = _fresh(blob, seed=8)
    o.forward()
    n0 = o.dim("nefc")
    hi = arrays["jnt_range"][0][1]
    o.qpos[0] = hi + 0.01
    o.forward()
    assert o.dim("nefc") == n0 + 1
    assert o.qfrc_constraint[0] < 0          # pushes back inside


def test_osc_zero_error_returns_bias(model):
    """osc.py:180-238 with x* = x, q* = q, zero velocities and q0 = q_arm: tau == qfrc_bias[arm] exactly."""
    blob = model[0]
    o, *_ = _fresh(blob, seed=9, jitter=0.4)
    o.forward()
    x = o.site_xpos[3].copy()
    R = o.site_xmat[3].reshape(3, 3)
    # quaternion of R via the oracle's own convergence check path: use scipy for independence
    from scipy.spatial.transform import Rotation
    qx = Rotation.from_matrix(R).as_quat()          # xyzw
    quat = np.array([qx[3], qx[0], qx[1], qx[2]])
    o.osc_config([350, 20, 500, 100, 200, 30], o.qpos[:7].copy())
    tg = np.zeros(13); tg[:3] = x; tg[6:10] = quat
    tau = o.osc_torque(tg)
    assert np.abs(tau - o.qfrc_bias[:7]).max() < 1e-9
    assert o.osc_converged(tg)


def test_osc_matches_numpy_restatement(model):
    """Line-by-line numpy restatement of mujoco_controllers/osc.py:99-238 on the oracle's J (finite differences), M."""
    blob = model[0]
    o, *_ = _fresh(blob, seed=10, jitter=0.4)
    o.qvel[:7] = [0.1, -0.2, 0.05, 0.3, -0.1, 0.2, -0.3]
    o.forward()
    nb = o.nbody
    x0 = o.site_xpos[3].copy(); R0 = o.site_xmat[3].reshape(3, 3).copy()
    eps = 1e-6
    J = np.zeros((6, 7))
    for d in range(7):
        q0 = o.qpos[d]
        o.qpos[d] = q0 + eps; o.forward(); xp, Rp = o.site_xpos[3].copy(), o.site_xmat[3].reshape(3, 3).copy()
        o.qpos[d] = q0 - eps; o.forward(); xm, Rm = o.site_xpos[3].copy(), o.site_xmat[3].reshape(3, 3).copy()
        o.qpos[d] = q0
        J[:3, d] = (xp - xm) / (2 * eps)
        W = (Rp - Rm) / (2 * eps) @ R0.T
        J[3:, d] = (W[2, 1], W[0, 2], W[1, 0])
    o.forward()
    Ma = o.qm[:7, :7]; Mi = np.linalg.inv(Ma)
    Lxi = J @ Mi @ J.T
    Lx = np.linalg.inv(Lxi) if abs(np.linalg.det(Lxi)) >= 1e-2 else np.linalg.pinv(Lxi, rcond=1e-2)
    from scipy.spatial.transform import Rotation
    qx = Rotation.from_matrix(R0).as_quat(); q = np.array([qx[3], qx[0], qx[1], qx[2]])
    tg = np.zeros(13); tg[:3] = x0 + (0.02, -0.1, 0.03); tg[6:10] = sc.DOWN_QUAT; tg[3:6] = (0.01, 0, 0)
    gains = [350, 20, 500, 100, 200, 30]
    q0n = sc.ARM_HOME
    o.osc_config(gains, q0n)

    def qmul(a, b):
        return np.array([a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3], a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
                         a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1], a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]])
    dx = J @ o.qvel[:7]
    pd_pos = gains[0] * np.clip(tg[:3] - x0, -0.05, 0.05) + gains[1] * (tg[3:6] - dx[:3])
    qe = qmul(tg[6:10] / np.linalg.norm(tg[6:10]), q * [1, -1, -1, -1])
    # mju_mat2Quat and scipy may differ by a global sign of q; sign(e_w) * e_xyz is invariant to it
    eo = qe[1:] * np.sign(qe[0])
    pd_ori = gains[2] * eo + gains[3] * (tg[10:13] - dx[3:])
    tau = J.T @ Lx @ np.concatenate([pd_pos, pd_ori])
    Nb = Mi @ J.T @ Lx
    tau += (np.eye(7) - J.T @ Nb.T) @ (gains[4] * (q0n - o.qpos[:7]) + gains[5] * (0 - o.qvel[:7]))
    tau += o.qfrc_bias[:7]
    got = o.osc_torque(tg)
    assert np.abs(got - tau).max() < 1e-5 * max(1, np.abs(tau).max())
    # null-space torque is dynamically consistent: J M^-1 (I - J' Nb') = 0
    assert np.abs(J @ Mi @ (np.eye(7) - J.T @ Nb.T)).max() < 1e-8


def test_halton_rollout_raises_no_flags(model):
    """Pattern of mujoco_menagerie/test/model_test.py:57-76: 1 s under Halton control noise, state stays finite."""
    blob = model[0]
    o, st, size, active = _fresh(blob, seed=11)
    o.set_option("full_contacts", 0)
    o.step1(); bias = o.qfrc_bias[:7].copy()
    for k in range(1000):
        o.ctrl[:] = sc.halton_ctrl(1, k, bias)[0]
        o.step()
    assert np.isfinite(o.qpos).all() and np.isfinite(o.qvel).all() and np.isfinite(o.qacc).all()
    assert np.abs(o.qvel).max() < 50


# ------------------------------------------------------------------------------------------ closed-form pins (SURVEY 8c fixture 2)
def closed_form_penetration(m=0.1, n_contacts=4, g=G, solref=(0.02, 1.0), solimp=(0.9, 0.95, 0.001, 0.5, 2.0), invweight=None):
    """Steady-state penetration r of a body of mass m resting on n contacts with MuJoCo's soft-constraint model (App. B item 9):
    each normal row carries f = D K imp r with D = 1/R, R = (1 - imp)/imp * diagApprox, diagApprox = 1/m (the table is static),
    K = 1/(dmax^2 timeconst^2 dampratio^2), imp = imp(r) from solimp  =>  r = m g (1 - imp) diagApprox / (n K imp^2), a fixed point."""
    d0, d1, w, mid, pw = solimp
    K = 1 / (d1 * d1 * solref[0] ** 2 * solref[1] ** 2)
    diag = (1 / m) if invweight is None else invweight
    r = 1e-4
    for _ in range(200):
        x = abs(r) / w
        y = 1.0 if x >= 1 else (x if pw == 1 else (x ** pw / mid ** (pw - 1) if x <= mid else 1 - (1 - x) ** pw / (1 - mid) ** (pw - 1)))
        imp = d0 + y * (d1 - d0)
        r = m * g * (1 - imp) * diag / (n_contacts * K * imp * imp)
    return r


def test_steady_state_penetration_is_the_solref_solimp_closed_form(model):
    """Pins K, imp(r), R = (1 - imp)/imp * diagApprox and geom_invweight0 jointly: the resting depth of the cube equals the fixed point
    of r = m g (1 - imp(r)) / (m 4 K imp(r)^2) with the default solref (0.02, 1) / solimp (0.9, 0.95, 0.001, 0.5, 2)."""
    blob = model[0]
    o, st, size, active = _fresh(blob, seed=2)
    o.set_option("full_contacts", 0)
    for _ in range(3000):
        o.step_frozen_arm()
    want = closed_form_penetration()
    assert 1.0e-4 < want < 1.2e-4
    for p in range(int(active.sum())):
        pen = 0.4 + size[p][2] - o.qpos[15 + 7 * p + 2]
        assert abs(pen - want) < 1e-9 * want + 1e-12, (pen, want)


def _incline(blob, tan_theta, mu, steps=400):
    import blobtools as bt
    th = np.arctan(tan_theta)
    o = Oracle(bt.tilted(blob, th, mu))
    st, size, active = sc.make_states(1, seed=2)
    active[0, 1:] = 0
    st[0, 15:18] = (0.45, 0.0, 0.4 + size[0, 0, 2] - 1e-4); st[0, 18:22] = (1, 0, 0, 0)
    sc.oracle_load(o, st[0], size[0], active[0]); o.set_option("full_contacts", 0)
    for _ in range(steps):
        o.step_frozen_arm()
    return o.qpos[15] - 0.45, o.qvel[15], th


def test_incline_sticks_iff_tan_theta_below_mu(model):
    """Cube on the table under gravity tilted by theta (mu = 0.5 < 1 so that it slides before it tips): it stays put for
    tan(theta) < mu (soft friction rows: a creep of ~0.1 mm/s, no more) and slides with a = g (sin - mu cos) for tan(theta) > mu."""
    blob = model[0]
    t = 0.4
    for tan_theta in (0.364, 0.45):
        x, v, th = _incline(blob, tan_theta, 0.5)
        assert abs(x) < 2e-4 and abs(v) < 1e-3, (tan_theta, x, v)
    for tan_theta in (0.55, 0.7):
        x, v, th = _incline(blob, tan_theta, 0.5)
        a = G * (np.sin(th) - 0.5 * np.cos(th))
        assert abs(x - 0.5 * a * t * t) < 0.05 * 0.5 * a * t * t, (tan_theta, x, 0.5 * a * t * t)


def pendulum_setup(blob, dof=1):
    """One-dof physical pendulum: the undamped arm with every dof but `dof` locked by a 1e8 armature.  -> (blob, q*, T) with q* the
    stable equilibrium and T = 2 pi sqrt(I / kappa), I = qM[dof, dof] (pinned by test_mass_matrix_from_independent_jacobians),
    kappa = d2 PE / dq2 from finite differences of sum_b m_b g z_com,b."""
    import blobtools as bt
    b = bt.free_arm(blob, lock_all_but=dof)
    counts, opts, arrays = B.unpack(b)
    mass = arrays["body_mass"]
    o = Oracle(b)
    st, size, active = sc.make_states(1, seed=2)
    active[0, :] = 0; st[0, :15] = 0

    def pe(q):
        s = st[0].copy(); s[dof] = q
        sc.oracle_load(o, s, size[0], active[0]); o.forward()
        return float((mass * G * o.xipos[:, 2]).sum())
    qs = np.linspace(-np.pi, np.pi, 721)
    q0 = qs[int(np.argmin([pe(q) for q in qs]))]
    eps, kappa = 1e-4, 0.0
    for _ in range(30):
        ep, em, e0 = pe(q0 + eps), pe(q0 - eps), pe(q0)
        kappa = (ep - 2 * e0 + em) / eps ** 2
        q0 -= (ep - em) / (2 * eps) / kappa
    pe(q0)
    inertia = o.qm[dof, dof]
    return b, st[0], size[0], active[0], q0, 2 * np.pi * np.sqrt(inertia / kappa)


def measure_period(step, get_q, q0, T, h=H):
    """mean time between downward zero crossings of q - q0 over ~3 periods (linear interpolation inside the step)"""
    ts, prev, t = [], get_q() - q0, 0.0
    for _ in range(int(3.3 * T / h)):
        step(); t += h
        cur = get_q() - q0
        if prev > 0 and cur <= 0:
            ts.append(t - h + h * prev / (prev - cur))
        prev = cur
    return float(np.mean(np.diff(ts)))


def test_pendulum_period(model):
    """Small oscillation (0.02 rad) of the locked-up arm about joint2: period = 2 pi sqrt(I / kappa) (1 + a^2/16), to 1e-5."""
    b, st, size, active, q0, T = pendulum_setup(model[0])
    o = Oracle(b)
    amp = 0.02
    s = st.copy(); s[1] = q0 + amp
    sc.oracle_load(o, s, size, active)
    got = measure_period(o.step, lambda: o.qpos[1], q0, T)
    assert abs(got / (T * (1 + amp * amp / 16)) - 1) < 1e-5, (got, T)


def _free_arm_energy(blob, h, integrator, T=0.5):
    import blobtools as bt
    b = bt.free_arm(blob, timestep=h, integrator=integrator)
    counts, opts, arrays = B.unpack(b)
    mass = arrays["body_mass"]
    o = Oracle(b)
    st, size, active = sc.make_states(1, seed=2)
    active[0, :] = 0
    st[0, :7] = sc.ARM_HOME; st[0, 7:15] = 0
    st[0, sc.REC_QVEL:sc.REC_QVEL + 7] = [0.5, -0.4, 0.3, 0.6, -0.5, 0.4, 0.3]
    sc.oracle_load(o, st[0], size[0], active[0])
    E = []
    for _ in range(int(round(T / h)) + 1):
        o.forward()
        v = o.qvel[:15]
        E.append(0.5 * v @ o.qm[:15, :15] @ v + float((mass * G * o.xipos[:, 2]).sum()))
        o.step()
    E = np.array(E)
    return float(np.abs(E - E[0]).max()), float(0.5 * v @ o.qm[:15, :15] @ v)


@pytest.mark.parametrize("integrator", [B.INT_IMPLICITFAST, B.INT_EULER])
def test_energy_drift_of_the_undamped_arm_is_first_order_in_h(model, integrator):
    """No damping, springs, limits, equalities, contacts or actuation: E = 1/2 v'Mv + sum m g z is conserved by the continuous system;
    the semi-implicit integrator's error is O(h) -- halving h halves the worst deviation (both integrators reduce to the same scheme)."""
    d1, ke = _free_arm_energy(model[0], 1e-3, integrator)
    d2, _ = _free_arm_energy(model[0], 5e-4, integrator)
    assert d1 < 0.01 * ke
    assert 1.8 < d1 / d2 < 2.2, (d1, d2)


@pytest.mark.parametrize("integrator", [B.INT_IMPLICITFAST, B.INT_EULER])
def test_integrator_matrix_against_numpy(model, integrator):
    """mj_Euler: (M + h diag(damping)) qacc+ = qfrc_smooth + qfrc_constraint; implicitfast additionally subtracts
    h * d(qfrc_actuator)/d(qvel) = -h * bias_vel * moment' moment of the affine gripper actuator wherever (i, j) lies in M's
    sparsity pattern (App. B item 15).  qvel' = qvel + h qacc+, recomputed with numpy from the oracle's M and forces."""
    import blobtools as bt
    blob, counts, opts, arrays = model
    b = bt.surgery(blob, opts=dict(integrator=integrator))
    o, st, size, active = _fresh(b, seed=5)
    o.qvel[:7] = np.linspace(-0.3, 0.3, 7); o.qvel[7] = 0.4; o.qvel[11] = 0.3
    o.ctrl[:] = [3, -2, 1, 1, 0.5, -0.5, 0.2, 100]
    o.set_option("full_contacts", 0)
    o.forward()
    nv = 15
    M = o.qm[:nv, :nv].copy(); f = (o.qfrc_smooth + o.qfrc_constraint)[:nv].copy(); v0 = o.qvel[:nv].copy()
    A = M + H * np.diag(arrays["dof_damping"][:nv])
    if integrator == B.INT_IMPLICITFAST:
        for a in range(counts["nu"]):
            bv = arrays["act_bias"][a][2]
            if bv == 0:
                continue
            for x in range(arrays["act_trnnum"][a]):
                for y in range(arrays["act_trnnum"][a]):
                    i, j = arrays["act_trndof"][a][x], arrays["act_trndof"][a][y]
                    if M[i, j] != 0 or i == j:
                        A[i, j] -= H * arrays["act_trncoef"][a][x] * arrays["act_trncoef"][a][y] * bv
    want = v0 + H * np.linalg.solve(A, f)
    o.step2()
    assert np.abs(o.qvel[:nv] - want).max() < 1e-9 * max(1.0, np.abs(want).max())
    # the two integrators really differ on this state (the drivers accelerate)
    if integrator == B.INT_EULER:
        o2, *_ = _fresh(blob, seed=5)
        o2.qvel[:7] = np.linspace(-0.3, 0.3, 7); o2.qvel[7] = 0.4; o2.qvel[11] = 0.3
        o2.ctrl[:] = [3, -2, 1, 1, 0.5, -0.5, 0.2, 100]; o2.set_option("full_contacts", 0); o2.step()
        assert np.abs(o2.qvel[:nv] - o.qvel[:nv]).max() > 1e-4
