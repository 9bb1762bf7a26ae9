"""CPU suite: the kernel's per-env templates compiled for the host (one-thread block; tests/hostemu) against the
oracle, on the same seeded scenes the GPU parity tests use.  This checks kernel LOGIC without a GPU; it is not a
product path (the package never loads the host build)."""
import numpy as np
import pytest

import scenarios as sc
from hostemu import MODE_FORWARD, MODE_OSC, MODE_PHYSICS, MODE_SETTLE, HostEmu, out_slices
from orc import Oracle

S = out_slices()


@pytest.fixture(scope="module")
def blob():
    return sc.default_blob()


@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-9), (np.float32, 2e-4)])
def test_forward_and_contact_sets(blob, dtype, tol):
    n = 6
    st, size, active = sc.make_states(n, seed=7)
    e = HostEmu(blob, dtype); o = Oracle(blob)
    r = e.run(MODE_FORWARD, st.astype(dtype), prop_size=size, prop_active=active, full_contacts=True)
    assert r["status"].max() == 0
    for i in range(n):
        out = r["out"][i].astype(np.float64)
        sc.oracle_load(o, st[i], size[i], active[i]); o.set_option("full_contacts", 1); o.forward()
        assert np.abs(out[S["site_xpos"]][:3] - o.site_xpos[3]).max() < tol
        assert np.abs(out[S["qfrc_bias"]][:sc.NV] - o.qfrc_bias).max() < tol * max(1, np.abs(o.qfrc_bias).max())
        assert np.abs(out[S["qacc"]][:sc.NV] - o.qacc).max() < 50 * tol * max(10, np.abs(o.qacc).max())
        oc = o.contacts(); nc = int(out[S["ncon"]][0])
        c = out[S["contact"]].reshape(-1, 10)[:nc]
        assert nc == len(oc["dist"])
        assert (c[:, 7].astype(int) == oc["geom"][:, 0]).all() and (c[:, 8].astype(int) == oc["geom"][:, 1]).all()
        assert (c[:, 9].astype(int) == (oc["exclude"] != 0)).all()
        assert np.abs(c[:, 0] - oc["dist"]).max(initial=0) < 10 * max(tol, 1e-6)


@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-8), (np.float32, 5e-3)])
def test_physics_steps(blob, dtype, tol):
    n, steps = 3, 100
    st, size, active = sc.make_states(n, seed=11, settled=False)
    e = HostEmu(blob, dtype); o = Oracle(blob)
    sc.oracle_load(o, st[0], size[0], active[0]); o.step1(); bias = o.qfrc_bias[:7].copy()
    ctrl = sc.halton_ctrl(n, 0, bias)
    got = st.astype(dtype)
    r = e.run(MODE_PHYSICS, got, n_steps=steps, ctrl=ctrl, prop_size=size, prop_active=active)
    assert r["status"].max() == 0
    for i in range(n):
        sc.oracle_load(o, st[i], size[i], active[i]); o.set_option("full_contacts", 0); o.ctrl[:] = ctrl[i]
        for _ in range(steps):
            o.step()
        want = sc.oracle_state(o)
        assert np.abs(got[i, :sc.NQ] - want[:sc.NQ]).max() < tol
        assert np.abs(got[i, sc.REC_QVEL:sc.REC_QVEL + sc.NV] - want[sc.REC_QVEL:sc.REC_QVEL + sc.NV]).max() < 100 * tol


@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-8), (np.float32, 5e-3)])
def test_osc_run_controller(blob, dtype, tol):
    n, ticks = 3, 40
    st, size, active = sc.make_states(n, seed=5)
    tg, grip = sc.random_targets(n, seed=3)
    e = HostEmu(blob, dtype); o = Oracle(blob)
    got = st.astype(dtype)
    r = e.run(MODE_OSC, got, n_steps=ticks, target=tg, grip=grip, prop_size=size, prop_active=active)
    for i in range(n):
        sc.oracle_load(o, st[i], size[i], active[i]); o.set_option("full_contacts", 0)
        c = o.run_controller(tg[i], int(grip[i]), ticks)
        assert np.abs(got[i, :sc.NQ] - o.qpos[:sc.NQ]).max() < tol
        assert bool(r["converged"][i]) == c


def test_settle(blob):
    st, size, active = sc.make_states(2, seed=21, settled=False)
    e = HostEmu(blob); o = Oracle(blob)
    got = st.copy()
    r = e.run(MODE_SETTLE, got, n_steps=2000, prop_size=size, prop_active=active, settle=(1e-3, 1e-2, 300))
    assert np.abs(got[:, :15] - st[:, :15]).max() == 0
    sc.oracle_load(o, st[0], size[0], active[0]); o.set_option("full_contacts", 0)
    k = 0
    act = np.repeat(active[0].astype(bool), 6)
    while k < 2000:
        o.step_frozen_arm(); k += 1
        if np.abs(o.qvel[15:39][act]).max() < 1e-3 and np.abs(o.qacc[15:39][act]).max() < 1e-2 and k > 300:
            break
    assert 300 < k < 2000 and k == r["steps_done"][0]
    assert np.abs(got[0, :sc.NQ] - o.qpos[:sc.NQ]).max() < 1e-9


@pytest.mark.parametrize("kind", ["grasp", "stack", "grasp_stack"])
@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-7), (np.float32, 5e-3)])
def test_coupled_blocks(blob, kind, dtype, tol):
    """Gripper squeezing a prop / stacked props: constraint rows couple the tree and prop blocks of the Newton Hessian
    (block-sparse Cholesky with fill-in), checked step by step against the oracle's dense solve."""
    st, size, active = sc.coupled_states(kind)
    e = HostEmu(blob, dtype); o = Oracle(blob)
    sc.oracle_load(o, st[0], size[0], active[0]); o.set_option("full_contacts", 0)
    o.step1(); ctrl = np.zeros((1, sc.NU)); ctrl[0, :7] = o.qfrc_bias[:7]; ctrl[0, 7] = 255.0
    got = st.astype(dtype); steps = 40
    r = e.run(MODE_PHYSICS, got, n_steps=steps, ctrl=ctrl, prop_size=size, prop_active=active)
    assert r["status"][0] == 0
    o.ctrl[:] = ctrl[0]; ncoupled = 0
    for _ in range(steps):
        o.step()
        ncoupled = max(ncoupled, o.dim("nefc"))
    assert ncoupled > 7 + 12                      # more than equality + one resting prop: pads / stacked faces touch
    assert np.abs(got[0, :sc.NQ] - o.qpos[:sc.NQ]).max() < tol
    assert np.abs(got[0, sc.REC_QVEL:sc.REC_QVEL + sc.NV] - o.qvel[:sc.NV]).max() < 200 * tol


def test_per_scene_gains(blob):
    """One tuner candidate per scene (b2p_set_env_gains): every scene follows the oracle configured with ITS gain row."""
    n, ticks = 3, 30
    st, size, active = sc.make_states(n, seed=8)
    tg, grip = sc.random_targets(n, seed=9)
    gains = np.array([[350, 20, 500, 100, 200, 30], [120, 35, 300, 60, 80, 12], [600, 50, 200, 20, 10, 5]], np.float64)
    e = HostEmu(blob, np.float64); o = Oracle(blob)
    got = st.copy()
    e.run(MODE_OSC, got, n_steps=ticks, target=tg, grip=grip, prop_size=size, prop_active=active, gains=gains)
    for i in range(n):
        sc.oracle_load(o, st[i], size[i], active[i]); o.set_option("full_contacts", 0)
        o.osc_config(gains[i], sc.ARM_HOME)
        o.run_controller(tg[i], int(grip[i]), ticks)
        assert np.abs(got[i, :sc.NQ] - o.qpos[:sc.NQ]).max() < 1e-8, i
    assert np.abs(got[0, :7] - got[1, :7]).max() > 1e-4        # and the rows really differ


def test_euler_integrator(blob):
    """SURVEY R6.12 / mj_Euler (the MEN-A menagerie snapshot's integrator family): the same scene compiled with integrator = Euler --
    joint damping alone enters the integrator matrix, the gripper actuator's velocity term does not (b2p_model.h impl_hD) -- against
    the oracle's mj_Euler branch, and it must differ from the implicitfast trajectory (else the test would not see the branch)."""
    import blobtools as bt
    from mujoco_robot_environments_b200.compile import blob as B
    eb = bt.surgery(blob, opts=dict(integrator=B.INT_EULER))
    n, steps = 2, 60
    st, size, active = sc.make_states(n, seed=13, settled=False)
    o = Oracle(eb)
    sc.oracle_load(o, st[0], size[0], active[0]); o.step1(); bias = o.qfrc_bias[:7].copy()
    ctrl = sc.halton_ctrl(n, 0, bias); ctrl[:, 7] = 200.0            # gripper closing: the actuator's -10 * velocity term is live
    got = st.copy()
    r = HostEmu(eb, np.float64).run(MODE_PHYSICS, got, n_steps=steps, ctrl=ctrl, prop_size=size, prop_active=active)
    ref = st.copy()
    HostEmu(blob, np.float64).run(MODE_PHYSICS, ref, n_steps=steps, ctrl=ctrl, prop_size=size, prop_active=active)
    assert r["status"].max() == 0
    for i in range(n):
        sc.oracle_load(o, st[i], size[i], active[i]); o.set_option("full_contacts", 0); o.ctrl[:] = ctrl[i]
        for _ in range(steps):
            o.step()
        want = sc.oracle_state(o)
        assert np.abs(got[i, :sc.NQ] - want[:sc.NQ]).max() < 1e-8
        assert np.abs(got[i, sc.REC_QVEL:sc.REC_QVEL + sc.NV] - want[sc.REC_QVEL:sc.REC_QVEL + sc.NV]).max() < 1e-6
    assert np.abs(got[:, :sc.NQ] - ref[:, :sc.NQ]).max() > 1e-6


def test_plane_convex_contacts(blob):
    """R6.6 plane-convex: an arm driven past its joint limits to below the table edge touches the ground plane with hull vertices --
    same contact list (geoms, distances, points) as the oracle's plane_convex, and a step from there stays finite."""
    st, size, active = sc.make_states(1, seed=3)
    o = Oracle(blob); e = HostEmu(blob, np.float64)
    poses = np.array([[-2.97, 2.3, -0.27, -0.19, -0.11, 2.36, 0.7], [-2.57, 1.89, -0.03, -0.87, -0.47, 2.45, 0.83], [-2.8, 2.03, 0.05, -0.6, 0.63, 2.42, 1.3]])
    seen = 0
    for q7 in poses:
        q = st[0].copy(); q[:7] = q7
        sc.oracle_load(o, q, size[0], active[0]); o.set_option("full_contacts", 0); o.forward()
        oc = o.contacts()
        r = e.run(MODE_FORWARD, q[None].copy(), prop_size=size, prop_active=active, full_contacts=False)
        out = r["out"][0]; nc = int(out[S["ncon"]][0]); c = out[S["contact"]].reshape(-1, 10)[:nc]
        assert r["status"][0] == 0 and nc == len(oc["dist"]) and (c[:, 7].astype(int) == oc["geom"][:, 0]).all() and (c[:, 8].astype(int) == oc["geom"][:, 1]).all()
        assert np.abs(c[:, 0] - oc["dist"]).max() < 1e-12 and np.abs(c[:, 1:4] - oc["pos"]).max() < 1e-12
        ground = (oc["geom"][:, 0] == 0) & (oc["geom"][:, 1] > 1)
        seen += int(ground.sum())
        assert (np.abs(c[ground, 4:7] - np.array([0, 0, 1.0])).max() < 1e-15) if ground.any() else True
    assert seen >= 4
