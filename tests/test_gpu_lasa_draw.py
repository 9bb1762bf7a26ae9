"""GPU side of tests/test_pyramidal_cone.py: the pyramidal friction cone and the drawing task (tasks/lasa_draw.py) through the C-ABI,
against the oracle on the same seeded inputs."""
import numpy as np
import pytest

import blobtools as bt
import scenarios as sc
from orc import Oracle

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")
HOME = [0, -0.785, 0, -2.356, 0, 1.571, 0.785]
TOUCH = [-0.59998, 0.56896, 0.44407, -2.42968, -0.90239, 2.84152, -1.69566]      # tool piece resting on the table (eef z = 0.5)


@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-9), (torch.float32, 2e-4)])
def test_pyramidal_cone_steps_parity(dtype, tol):
    """The rearrangement scene with <option cone="pyramidal">: 250 steps under Halton torque noise, cubes dropped onto the table."""
    from mujoco_robot_environments_b200.physics import BatchedPhysics
    blob = bt.surgery(sc.default_blob(), opts={"cone": 0})
    n, steps = 16, 250
    st, size, active = sc.make_states(n, seed=5, settled=False)
    o = Oracle(blob)
    sc.oracle_load(o, st[0], size[0], active[0]); o.step1(); bias = o.qfrc_bias[:7].copy()
    ctrl = sc.halton_ctrl(n, 0, bias)
    p = BatchedPhysics(n, dtype=dtype, blob=blob)
    p.state.copy_(torch.tensor(st, dtype=dtype)); p.set_env_params(torch.tensor(size), torch.tensor(active))
    p.set_control(torch.tensor(ctrl)); p.step(steps); torch.cuda.synchronize()
    got = p.state.cpu().numpy().astype(np.float64)
    assert int(p.status().cpu().max()) == 0
    assert int(p.data.nefc.cpu().max()) >= 41                        # four edge rows per contact
    for e in range(n):
        sc.oracle_load(o, st[e], size[e], active[e]); o.set_option("full_contacts", 0)
        o.ctrl[:8] = ctrl[e]
        for _ in range(steps):
            o.step()
        assert np.abs(got[e, :sc.NQ] - o.qpos[:sc.NQ]).max() < tol, e
    with pytest.raises(RuntimeError):
        p.set_solver("pgs")                                           # the PGS instantiation covers the elliptic cone only


@pytest.mark.parametrize("dtype,ticks,tol", [(torch.float64, 20, 1e-9), (torch.float32, 8, 1e-4)])
def test_move_to_draw_target_parity(dtype, ticks, tol):
    from mujoco_robot_environments_b200.lasa_draw import BatchedLasaDrawEnv, TOOL_OFFSET
    from mujoco_robot_environments_b200.physics import load_default_blob
    n = 4
    env = BatchedLasaDrawEnv(n, dtype=dtype)
    o = Oracle(load_default_blob("lasa_draw")[0]); o.reset(); o.qpos[:7] = HOME
    home = env.reset()
    o.forward()
    assert np.abs(home[0].cpu().numpy() - o.site_xpos[-1]).max() < 1e-5
    for t in range(ticks):
        ang = 2 * np.pi * t / 100
        pos = np.array([0.40 + 0.05 * np.cos(ang), 0.05 * np.sin(ang), 0.7]); vel = np.array([-0.06 * np.sin(ang), 0.06 * np.cos(ang), 0])
        q, v, u = env.move_to_draw_target(torch.tensor(pos).repeat(n, 1), torch.tensor(vel).repeat(n, 1))
        tg = np.concatenate([pos + TOOL_OFFSET, vel, [0, 0, 1, 0], [0, 0, 0]])
        if t == ticks - 1:
            o.step1(); want_u = o.osc_torque(tg)
        o.run_controller(tg, 0, 1)
    torch.cuda.synchronize()
    assert q.shape == (n, 7) and v.shape == (n, 7) and u.shape == (n, 7)
    assert np.abs(q.cpu().numpy().astype(np.float64) - o.qpos[:7]).max() < tol
    assert torch.equal(q[0], q[n - 1])                                # identical scenes, identical results
    assert np.abs(u[0].cpu().numpy() - want_u).max() < max(1e3 * tol, 1e-6) * max(1.0, np.abs(want_u).max())   # the torque of the last tick
    assert int(env.physics.status().cpu().max()) == 0


@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-9), (torch.float32, 1e-4)])
def test_move_to_joint_position_target_with_the_tool_on_the_table(dtype, tol):
    from mujoco_robot_environments_b200.lasa_draw import BatchedLasaDrawEnv
    from mujoco_robot_environments_b200.physics import load_default_blob
    n = 3
    env = BatchedLasaDrawEnv(n, dtype=dtype, actuators="position")
    env.physics.set_out_fields("SITE", "DYN", "EFC")
    o = Oracle(load_default_blob("lasa_draw_position")[0]); o.reset(); o.qpos[:7] = TOUCH
    env.reset(TOUCH)
    nefc_max = 0
    for t in range(30):
        target = np.array(TOUCH) + np.array([0.004 * t, 0.002, 0, 0, 0, 0.001, 0])
        q = env.move_to_joint_position_target(torch.tensor(target))
        nefc_max = max(nefc_max, int(env.physics.data.nefc.cpu().max()))
        o.ctrl[:7] = target
        for _ in range(5):
            o.step()
    assert nefc_max >= 4                                              # the cylinder touched the table: four pyramid edges
    assert torch.equal(q[0], q[n - 1])                                # identical scenes, identical results (a 7-dof model once read shared memory past its rows)
    assert np.abs(q.cpu().numpy().astype(np.float64) - o.qpos[:7]).max() < tol
    with pytest.raises(RuntimeError):
        env.move_to_draw_target(torch.zeros(n, 3), torch.zeros(n, 3))


def test_main_camera_observation_bit_exact_vs_render_oracle():
    """lasa_draw.py:281-298 on the CUDA ray-caster: rgb bytes and depth floats of the drawing scene's own render model (tool piece as a
    64-gon prism, draw plane as a blended zone) against the plain-C restatement, on a window around the arm and on a full frame."""
    import rnd
    from mujoco_robot_environments_b200.lasa_draw import BatchedLasaDrawEnv
    from mujoco_robot_environments_b200.render import load_default_render_blob
    n = 2
    env = BatchedLasaDrawEnv(n, dtype=torch.float32)
    env.physics.qpos[1, :7] = torch.tensor(TOUCH, device=env.device)
    env.physics.forward(full_contacts=False)
    obs = env.step()
    torch.cuda.synchronize()
    assert obs["main_camera/rgb"].shape == (n, 640, 640, 3) and obs["main_camera/depth"].shape == (n, 640, 640)
    xp, xm = env.physics.data.geom_xpos.cpu().numpy().astype(np.float64), env.physics.data.geom_xmat.cpu().numpy().astype(np.float64)
    blob = load_default_render_blob("lasa_draw")
    for e, win in ((0, None), (1, (160, 96, 544, 480))):
        rgb, depth, seg = rnd.render(blob, xp[e], xm[e], None, None, None, window=win)
        x0, y0, x1, y1 = win if win else (0, 0, 640, 640)
        g_rgb, g_d = obs["main_camera/rgb"][e].cpu().numpy()[y0:y1, x0:x1], obs["main_camera/depth"][e].cpu().numpy()[y0:y1, x0:x1]
        assert np.array_equal(g_d.view(np.uint32), depth[y0:y1, x0:x1].view(np.uint32)), ("depth", e)
        assert np.array_equal(g_rgb, rgb[y0:y1, x0:x1]), ("rgb", e)
        assert (seg[y0:y1, x0:x1] == 1).any() and ((seg[y0:y1, x0:x1] >= 3) & (seg[y0:y1, x0:x1] <= 12)).any()      # table and arm in view
    assert not torch.equal(obs["main_camera/rgb"][0], obs["main_camera/rgb"][1])
