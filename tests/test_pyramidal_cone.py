"""The pyramidal friction cone (MuJoCo's default; the scenes without the gripper's XML -- lasa_draw / push -- run it) and the drawing
scene itself.  The oracle's pyramidal rows against analytic answers, its two solvers against each other, and the host build of the
kernel templates (class-capable instantiation, ROW_PYR rows) against the oracle.  The cone's formulae (edge Jacobians, diagApprox,
R_py = 2 mu_reg^2 R) are restated from MuJoCo's published algorithm: parity unpinned, like the rest of the oracle."""
import os

import numpy as np
import pytest

import blobtools as bt
import hostemu as he
import scenarios as sc
from hostemu import MODE_OSC, MODE_PHYSICS, HostEmu
from orc import Oracle

G = 9.8
ASSETS = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "mujoco_robot_environments_b200", "assets")
HOME = [0, -0.785, 0, -2.356, 0, 1.571, 0.785]
TOUCH = [-0.59998, 0.56896, 0.44407, -2.42968, -0.90239, 2.84152, -1.69566]      # tool piece resting on the table (eef z = 0.5)


@pytest.fixture(scope="module")
def pyr_blob():
    return bt.surgery(sc.default_blob(), opts={"cone": 0})           # the rearrangement scene with <option cone="pyramidal">


def _draw_blob(tag="lasa_draw"):
    with open(os.path.join(ASSETS, tag + ".b2pm"), "rb") as f:
        return f.read()


def test_resting_cube_weight_and_penetration(pyr_blob):
    """Settled cubes: the edge forces of a contact sum to the normal force, the normal forces to m g; the steady-state penetration
    follows from R_py: four edges of R_py = 2 (mu^2 / impratio) (1 + mu^2) R_n each are 10x as stiff as the elliptic normal row
    (mu = 1, impratio = 10), so the cube sinks a tenth as deep."""
    o = Oracle(pyr_blob); oe = Oracle(sc.default_blob())
    st, size, active = sc.make_states(1, seed=2)
    for q in (o, oe):
        sc.oracle_load(q, st[0], size[0], active[0]); q.set_option("full_contacts", 0)
        for _ in range(600):
            q.step_frozen_arm()
        q.forward()
    for p in range(int(active[0].sum())):
        assert abs(o.qfrc_constraint[15 + 6 * p + 2] - 0.1 * G) < 1e-6
        pen, pen_e = 0.4 + size[0][p][2] - o.qpos[15 + 7 * p + 2], 0.4 + size[0][p][2] - oe.qpos[15 + 7 * p + 2]
        assert 0 < pen < 5e-5
        assert abs(pen / pen_e - 0.1) < 0.02, (pen, pen_e)


def _incline(blob, tan_theta, mu, steps=400):
    th = np.arctan(tan_theta)
    o = Oracle(bt.surgery(bt.tilted(blob, th, mu), opts={"cone": 0}))
    st, size, active = sc.make_states(1, seed=2)
    active[0, 1:] = 0
    st[0, 15:18] = (0.45, 0.0, 0.4 + size[0, 0, 2] - 1e-5); st[0, 18:22] = (1, 0, 0, 0)
    sc.oracle_load(o, st[0], size[0], active[0]); o.set_option("full_contacts", 0)
    for _ in range(steps):
        o.step_frozen_arm()
    return o.qpos[15] - 0.45, o.qvel[15], th


def test_incline_sticks_iff_tan_theta_below_mu():
    """Along a pyramid axis the pyramidal cone's friction limit is mu f_n exactly: the cube stays for tan(theta) < mu and slides with
    a = g (sin - mu cos) beyond."""
    blob = sc.default_blob()
    for tan_theta in (0.364, 0.45):
        x, v, th = _incline(blob, tan_theta, 0.5)
        assert abs(x) < 1e-3 and abs(v) < 1e-2, (tan_theta, x, v)      # soft rows creep (a sliding cube reaches 0.17 m/s at tan = 0.55)
    for tan_theta in (0.55, 0.7):
        x, v, th = _incline(blob, tan_theta, 0.5)
        a = G * (np.sin(th) - 0.5 * np.cos(th))
        assert abs(x - 0.5 * a * 0.4 * 0.4) < 0.05 * 0.5 * a * 0.4 * 0.4, (tan_theta, x)


def test_newton_and_pgs_agree_on_pyramidal_rows(pyr_blob):
    """Two different minimisers of one convex problem (primal Newton, dual PGS): with a pyramidal cone every contact row is a scalar
    inequality row in both."""
    st, size, active = sc.make_states(2, seed=9, settled=False)
    for e in range(2):
        o = Oracle(pyr_blob)
        sc.oracle_load(o, st[e], size[e], active[e]); o.set_option("full_contacts", 0)
        for _ in range(80):
            o.step()                                   # falls 2 cm, lands
        state = sc.oracle_state(o)
        qa = []
        for solver, its, tol in ((0, 100, 1e-12), (1, 50000, 1e-17)):              # as tests/test_solver_crosscheck.py runs the two to convergence
            sc.oracle_load(o, state, size[e], active[e]); o.set_option("full_contacts", 0)
            o.set_option("solver", solver); o.set_option("iterations", its); o.set_option("tolerance", tol)
            o.forward()
            qa.append(np.array(o.qacc[:sc.NV]).copy())
        assert np.abs(qa[0]).max() > 1.0                                           # a live scene: the cubes have just landed
        assert np.abs(qa[0] - qa[1]).max() < 1e-5 * max(1.0, np.abs(qa[0]).max()), np.abs(qa[0] - qa[1]).max()


@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-11), (np.float32, 2e-4)])
def test_kernel_host_build_matches_oracle_on_pyramidal_scene(pyr_blob, dtype, tol):
    """250 steps under Halton torque noise, props dropped 2 cm onto the table (4 edge rows per contact, 41..73 rows per scene)."""
    n, steps = 6, 250
    st, size, active = sc.make_states(n, seed=5, settled=False)
    o = Oracle(pyr_blob)
    sc.oracle_load(o, st[0], size[0], active[0]); o.step1(); bias = o.qfrc_bias[:7].copy()
    ctrl = sc.halton_ctrl(n, 0, bias)
    h = HostEmu(pyr_blob, dtype)
    s2 = st.astype(dtype).copy()
    r = h.run(MODE_PHYSICS, s2, n_steps=steps, ctrl=ctrl.astype(dtype), prop_size=size.astype(dtype), prop_active=active)
    sl = he.out_slices()
    assert (r["status"] == 0).all() and r["out"][:, sl["nefc"]].max() >= 41
    for e in range(n):
        sc.oracle_load(o, st[e], size[e], active[e]); o.set_option("full_contacts", 0)
        o.ctrl[:8] = ctrl[e]
        for _ in range(steps):
            o.step()
        assert np.abs(s2[e, :sc.NQ].astype(np.float64) - o.qpos[:sc.NQ]).max() < tol


@pytest.mark.parametrize("dtype,ticks,tol", [(np.float64, 20, 1e-9), (np.float32, 8, 1e-4)])
def test_draw_scene_osc_ticks_match_oracle(dtype, ticks, tol):
    """lasa_draw.py: move_to_draw_target = one OSC tick (velocity target, 5 steps) on the arm-only scene, from home.  With
    config/lasa.yaml's 10 ms step the loop is not stable (joint speeds reach 17 rad/s), so rounding differences grow tick by tick:
    the fp32 comparison stops after 8 ticks (3e-5 there, 6e-3 after 20)."""
    blob = _draw_blob()
    o = Oracle(blob); o.reset(); o.qpos[:7] = HOME
    h = HostEmu(blob, dtype)
    st = np.zeros((1, sc.REC_SIZE), dtype); st[0, sc.REC_QPOS:sc.REC_QPOS + 7] = HOME
    for t in range(ticks):
        ang = 2 * np.pi * t / 100
        tg = np.concatenate([[0.40 + 0.05 * np.cos(ang), 0.05 * np.sin(ang), 0.8], [-0.06 * np.sin(ang), 0.06 * np.cos(ang), 0], [0, 0, 1, 0], [0, 0, 0]])
        o.run_controller(tg, 0, 1)
        r = h.run(MODE_OSC, st, n_steps=1, control_steps=5, target=tg[None].astype(dtype), grip=np.zeros(1, np.uint8))
        assert r["status"][0] == 0
    assert np.abs(st[0, sc.REC_QPOS:sc.REC_QPOS + 7].astype(np.float64) - o.qpos[:7]).max() < tol
    assert np.abs(o.qpos[:7] - HOME).max() > 0.05                    # the arm did move


@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-9), (np.float32, 1e-4)])
def test_draw_scene_tool_on_table_matches_oracle(dtype, tol):
    """lasa_draw.py: move_to_joint_position_target on the position-servo scene, the tool piece pressed on the table and dragged:
    cylinder (on the tree) against box through the convex query, one contact, four pyramid edges."""
    blob = _draw_blob("lasa_draw_position")
    o = Oracle(blob); o.reset(); o.qpos[:7] = TOUCH
    h = HostEmu(blob, dtype)
    st = np.zeros((1, sc.REC_SIZE), dtype); st[0, sc.REC_QPOS:sc.REC_QPOS + 7] = TOUCH
    sl = he.out_slices(); nefc_max = 0
    for t in range(30):
        target = np.array(TOUCH) + np.array([0.004 * t, 0.002, 0, 0, 0, 0.001, 0])        # swing the base joint, lean on the table
        o.ctrl[:7] = target
        for _ in range(5):
            o.step()
        r = h.run(MODE_PHYSICS, st, n_steps=5, ctrl=np.concatenate([target, [0]])[None].astype(dtype))
        assert r["status"][0] == 0
        nefc_max = max(nefc_max, int(r["out"][0, sl["nefc"]][0]))
    assert nefc_max >= 4                                              # the tool touched: four edge rows
    assert np.abs(st[0, sc.REC_QPOS:sc.REC_QPOS + 7].astype(np.float64) - o.qpos[:7]).max() < tol
