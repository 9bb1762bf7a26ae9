"""GPU parity of the CUDA ray-cast observations (csrc/b2p_render.cu) through the C-ABI (include/b2p_render.h) against the plain-C
restatement (oracle/b2p_render_oracle.c): BIT-EXACT rgb bytes, depth floats and segmentation ids -- the kernel is built without
FMA contraction for exactly this reason -- plus size-independent properties at batch sizes the oracle could not render."""
import numpy as np
import pytest

import rnd
import scenarios as sc

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")
REACH = [0.0, 0.35, 0.0, -1.9, 0.0, 2.25, 0.785]


def _setup(n, dtype, seed=11, reach_every=2):
    from mujoco_robot_environments_b200.physics import BatchedPhysics
    from mujoco_robot_environments_b200.render import BatchedRenderer
    st, size, active = sc.make_states(n, seed=seed)
    for e in range(0, n, reach_every):           # half of the scenes: arm stretched into the camera view
        st[e, :7] = np.asarray(REACH) + 0.1 * (np.random.default_rng(e).random(7) - 0.5)
    p = BatchedPhysics(n, dtype=dtype)
    p.state.copy_(torch.tensor(st, dtype=dtype)); p.set_env_params(torch.tensor(size), torch.tensor(active))
    p.forward(full_contacts=False)
    rgba = torch.tensor(np.random.default_rng(seed).uniform(-0.1, 1.1, (n, 4, 4)).astype(np.float32))
    return p, BatchedRenderer(p.device), rgba, size, active


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_render_bit_exact_vs_oracle(dtype):
    n = 3
    p, r, rgba, size, active = _setup(n, dtype)
    out = r.render(p, rgba)
    torch.cuda.synchronize()
    assert r.query("TILES_CAST") > 0
    xp, xm = p.data.geom_xpos.cpu().numpy().astype(np.float64), p.data.geom_xmat.cpu().numpy().astype(np.float64)
    blob = rnd.render_blob()
    for e in range(n):
        # scene 0 in full; the others on a window that covers the arm, the zones and most props (keeps the CPU side short)
        win = None if e == 0 else (64, 160, 576, 480)
        rgb, depth, seg = rnd.render(blob, xp[e], xm[e], rgba[e].numpy(), p.prop_size[e].cpu().numpy().astype(np.float64), active[e], window=win)
        x0, y0, x1, y1 = win if win else (0, 0, 640, 480)
        g_rgb, g_d, g_s = (out[k][e].cpu().numpy()[y0:y1, x0:x1] for k in ("rgb", "depth", "seg"))
        assert np.array_equal(g_s, seg[y0:y1, x0:x1]), ("seg", e, int((g_s != seg[y0:y1, x0:x1]).sum()))
        assert np.array_equal(g_d.view(np.uint32), depth[y0:y1, x0:x1].view(np.uint32)), ("depth", e)
        assert np.array_equal(g_rgb, rgb[y0:y1, x0:x1]), ("rgb", e)
        assert ((g_s >= 2) & (g_s <= 25)).any() == (e % 2 == 0)


def test_render_batch_properties():
    """300 scenes: identical scenes give identical frames wherever they sit in the batch; most tiles take the background-copy
    path; scenes with the arm at home equal the static background outside the prop tiles; outputs may be requested separately."""
    n = 300
    p, r, rgba, size, active = _setup(n, torch.float32, seed=3, reach_every=3)
    p.state[150:] = p.state[:150]; p.prop_size[150:] = p.prop_size[:150]; p.prop_active[150:] = p.prop_active[:150]; rgba[150:] = rgba[:150]
    p.forward(full_contacts=False)
    out = r.render(p, rgba)
    torch.cuda.synchronize()
    for k in ("rgb", "depth", "seg"):
        assert torch.equal(out[k][:150], out[k][150:]), k
    copied, cast = r.query("TILES_COPIED"), r.query("TILES_CAST")
    assert copied + cast == n * 1200 and copied > 0.6 * n * 1200, (copied, cast)
    # a home-pose scene: every pixel is table / zone / prop (no arm, no empty pixel), depth within [table - prop height, table]
    seg, d = out["seg"][1], out["depth"][1]
    assert int(((seg >= 2) & (seg <= 25)).sum()) == 0 and int((seg < 0).sum()) == 0
    assert float(d.max()) <= 0.9 + 1e-6 and float(d.min()) > 0.9 - 0.04
    # depth-only request equals the full request
    only = {"rgb": out["rgb"].clone(), "depth": torch.zeros_like(out["depth"])}
    r.render(p, rgba, out=only, seg=False)
    assert torch.equal(only["depth"], out["depth"]) and torch.equal(only["rgb"], out["rgb"])


def test_env_render_and_pixel_2_world():
    from mujoco_robot_environments_b200 import config as C
    from mujoco_robot_environments_b200.env import BatchedRearrangementEnv
    from mujoco_robot_environments_b200.data_generation import COLOUR_SPLITTER_OVERRIDES
    env = BatchedRearrangementEnv(cfg=C.compose(COLOUR_SPLITTER_OVERRIDES), num_envs=8, strict=False)
    ts = env.reset()
    assert "overhead_camera/rgb" not in ts.observation          # batch default: state observation only
    obs = env.render(seg=True)
    assert obs["overhead_camera/rgb"].shape == (8, 480, 640, 3) and obs["overhead_camera/rgb"].dtype == torch.uint8
    assert obs["overhead_camera/depth"].shape == (8, 480, 640) and obs["overhead_camera/depth"].dtype == torch.float32
    pos, _ = env.expert.prop_geom_pose()
    top = pos[:, 0].clone(); top[:, 2] += env.prop_half_size[:, 0, 2]
    px = env.world_2_pixel("overhead_camera/overhead_camera", top)
    e = torch.arange(8, device=env.device)
    assert bool((obs["segmentation"][e, px[:, 1].long(), px[:, 0].long()] == int(env._prop_gid[0])).all())
    # rendered colour of prop 0 follows its sampled rgba (dominant channel of the label colour stays dominant)
    c = obs["overhead_camera/rgb"][e, px[:, 1].long(), px[:, 0].long()].float()
    base = env.prop_rgba[:, 0, :3]
    assert bool((c.argmax(1) == base.argmax(1)).all() | (base.max(1).values - base.min(1).values < 0.3).all())
    # bounding boxes from the segmentation image contain the projected prop centre and are about one prop wide
    bb = env.prop_bboxes()
    b0 = bb[:, 0]
    assert bool(((b0[:, 0] <= px[:, 0]) & (px[:, 0] <= b0[:, 2]) & (b0[:, 1] <= px[:, 1]) & (px[:, 1] <= b0[:, 3])).all())
    assert bool(((b0[:, 2] - b0[:, 0]) < 30).all() & ((b0[:, 2] - b0[:, 0]) >= 12).all())
    assert bool((bb[~env.has_prop] == -1).all())
    w = env.pixel_2_world("overhead_camera/overhead_camera", px.to(torch.float64))
    assert float((w - top.double()).abs().max()) < 3e-3


def test_single_scene_facade_observation_is_the_camera_pair():
    """RearrangementEnv(viewer=None, cfg=...) keeps the reference's observation keys and specs (rearrangement.py:442-478)."""
    from mujoco_robot_environments_b200.env import RearrangementEnv
    env = RearrangementEnv(dtype=torch.float64)
    ts = env.reset()
    o = ts.observation
    assert o["overhead_camera/rgb"].shape == (1, 480, 640, 3) and o["overhead_camera/depth"].shape == (1, 480, 640)
    spec = env.observation_spec()
    assert spec["overhead_camera/rgb"] == (1, 480, 640, 3) and spec["overhead_camera/depth"] == (1, 480, 640)
    assert float(o["overhead_camera/depth"].max()) <= 0.9 + 1e-6
