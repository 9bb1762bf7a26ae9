"""Scripted-expert maths against the reference's own formulae evaluated with numpy / scipy, on CPU; the batched
episode itself needs the kernels and is a gpu test."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
from scipy.spatial.transform import Rotation as R

from mujoco_robot_environments_b200 import config as C
from mujoco_robot_environments_b200 import expert as X
from mujoco_robot_environments_b200.data_generation import COLOUR_SPLITTER_OVERRIDES, pack_action


def test_grasp_quat_matches_reference_formula():
    """prop_pick (rearrangement.py:585-593) restated with scipy exactly as the reference calls it."""
    rng = np.random.default_rng(0)
    for _ in range(200):
        m = R.from_euler("z", rng.uniform(-np.pi, np.pi)).as_matrix()
        z = abs(np.rad2deg(np.arctan2(m[1, 0], m[0, 0]))); z = min([z, z - 90])
        ref = R.from_euler("xyz", [0, 180, z], degrees=True).as_matrix()
        q = X.grasp_quat_from_xmat(torch.tensor(m.reshape(1, 9)))[0].numpy()
        assert abs(np.linalg.norm(q) - 1) < 1e-12
        assert np.abs(R.from_quat([q[1], q[2], q[3], q[0]]).as_matrix() - ref).max() < 1e-12


def test_euler_xyz_is_scipy_extrinsic():
    rng = np.random.default_rng(1)
    a = rng.uniform(-3, 3, (50, 3))
    got = X.euler_xyz_matrix(*[torch.tensor(a[:, k]) for k in range(3)]).numpy().reshape(-1, 3, 3)
    assert np.abs(got - R.from_euler("xyz", a).as_matrix()).max() < 1e-12


def test_zone_table_and_colour_sampling():
    cfg = C.compose(COLOUR_SPLITTER_OVERRIDES)
    assert cfg["name"] == "colour_splitter"
    lo, hi = X.zone_table(cfg["task"])
    r = X.COLOUR_NAMES.index("red")       # red -> top_left: location (0.55, 0.4), size (0.075, 0.15) => +- size / 2
    assert np.allclose(lo[r], (0.5125, 0.325, 0.4)) and np.allclose(hi[r], (0.5875, 0.475, 0.4))
    idx, rgba = X.sample_prop_colours(500, 4, cfg["arena"]["props"]["colours"], np.random.default_rng(3))
    cols = cfg["arena"]["props"]["colours"]
    assert (idx[:, 0] == X.COLOUR_NAMES.index(cols[0])).all() and (idx[:, 1] == X.COLOUR_NAMES.index(cols[1])).all()   # props.py:607-610
    assert len(np.unique(idx[:, 2:])) == 6
    base = np.array([X.COLOURS[c] for c in X.COLOUR_NAMES])[idx]
    assert np.abs(rgba[..., :3] - base[..., :3]).max() <= 0.1 + 1e-6 and (rgba[..., 3] == 1).all()


def test_camera_maps_match_reference_formulae():
    """world_2_pixel / pixel_2_world (rearrangement.py:507-550) restated with numpy; round trip through a known depth."""
    cam = X.Camera(C.DEFAULT_CONFIG["arena"]["cameras"][0])
    K, E = cam.intrinsics(), cam.extrinsics()
    f = (1.0 / np.tan(np.deg2rad(61) / 2)) * 480 / 2.0
    assert np.allclose(K, [[-f, 0, 319.5], [0, f, 239.5], [0, 0, 1]])
    pts = np.random.default_rng(2).uniform([0.3, -0.4, 0.4], [0.6, 0.4, 0.45], (64, 3))
    px = cam.world_2_pixel(torch.tensor(pts)).numpy()
    for i, c in enumerate(pts):
        cc = E @ np.concatenate([c, np.ones(1)]); cc = cc[:3] / cc[3]
        ic = K @ cc; ic = ic[:2] / ic[2]
        assert (np.round(ic).astype(np.int32) == px[i]).all()
        assert 0 <= px[i, 0] < 640 and 0 <= px[i, 1] < 480            # the workspace is inside the overhead view
        depth = -cc[2]                                                  # distance along the viewing axis (camera looks down -z)
        w = cam.pixel_2_world(torch.tensor(ic[None]), torch.tensor([depth]))[0].numpy()
        assert np.abs(w - c).max() < 1e-9
    a = pack_action(torch.zeros(3, 7), torch.ones(3, 2, dtype=torch.int32))
    assert a.shape == (3, 10) and a.dtype == torch.float64


@pytest.mark.gpu
def test_sort_colours_episode_on_gpu():
    """Batched expert episode: pick poses sit on the props, place poses lie inside the colour's zone and are collision free
    by the reference's rule, the live state is untouched by the placement query, and after the episode most props are sorted."""
    from mujoco_robot_environments_b200.env import BatchedRearrangementEnv
    from mujoco_robot_environments_b200.data_generation import rollout_episode
    env = BatchedRearrangementEnv(cfg=C.compose(COLOUR_SPLITTER_OVERRIDES), num_envs=96, strict=False)
    env.reset()
    ex = env.expert
    before = env.physics.state.clone()
    n0 = int((env.has_prop & ex.within_target()).sum())
    ip, pick, place, pid = ex.sort_colours()
    assert torch.equal(before, env.physics.state)
    assert bool(ip.any())
    pos, mat = ex.prop_geom_pose()
    e = torch.arange(env.num_envs, device=env.device)
    assert torch.allclose(pick[ip, :3], pos[e, pid][ip])
    colour = env.prop_colour[e, pid]
    lo, hi = ex.zone_lo[colour], ex.zone_hi[colour]
    assert bool(((place[ip, :2] >= lo[ip, :2] - 1e-6) & (place[ip, :2] <= hi[ip, :2] + 1e-6)).all())
    # the reference's rule: no other prop closer than 0.05 (surface distance) -> centres further apart than 0.05
    others = pos.clone(); others[e, pid] = 100.0
    d = (others[:, :, :2] - place[:, None, :2]).norm(dim=2)
    assert bool((d[ip][env.has_prop[ip]] > 0.05).all())
    # independent check of the device-side search (b2p_place_search): put every placed prop at its pose, list the contacts with
    # b2p_forward and apply the reference's rule (rearrangement.py:617-625) on the host side
    p = env.physics
    saved, saved_out = p.state.clone(), p.out.clone()
    qadr = torch.tensor(p.model.prop_qposadr, device=env.device)[pid]
    cols = qadr[:, None] + torch.arange(7, device=env.device)[None, :]
    p.qpos[e[:, None], cols] = torch.where(ip[:, None], place.to(p.dtype), p.qpos[e[:, None], cols])
    p.forward(full_contacts=True)
    c, gid = p.data.contact, env._prop_gid[pid]
    touches = (c["geom1"] == gid[:, None]) | (c["geom2"] == gid[:, None])
    table = (c["geom1"] == env._table_gid) | (c["geom2"] == env._table_gid)
    slots = torch.arange(p.maxcon, device=env.device)[None, :]
    coll = ((slots < p.data.ncon[:, None]) & touches & ~table & (c["dist"] <= ex.PLACE_DIST)).any(1)
    assert not bool(coll[ip].any())
    p.state.copy_(saved); p.out.copy_(saved_out)
    assert ex.place_attempts >= 1
    out = rollout_episode(env, 4)
    n1 = int((env.has_prop & ex.within_target()).sum())
    total = int(env.has_prop.sum())
    assert out["actions"].shape == (96, 4, 2, 10) and out["valid"][:, 0].any()
    assert n1 > n0 and n1 > 0.6 * total, (n0, n1, total)
    assert int((env.physics.status() & 7).ne(0).sum()) == 0          # no BADQPOS / BADQVEL / BADQACC during the grasping episode


@pytest.mark.gpu
def test_recycled_rollout_on_gpu():
    """Finished scenes start their next episode while the others carry on: at least the requested number of episodes comes
    back in the record layout of rollout_episode, rounds never exceed max_steps, and a successful episode is a sorted scene."""
    from mujoco_robot_environments_b200.env import BatchedRearrangementEnv
    from mujoco_robot_environments_b200.data_generation import rollout_recycled
    env = BatchedRearrangementEnv(cfg=C.compose(COLOUR_SPLITTER_OVERRIDES), num_envs=48, strict=False)
    out = rollout_recycled(env, 80, 3)
    m = out["success"].shape[0]
    assert m >= 80
    assert out["actions"].shape == (m, 3, 2, 10) and out["valid"].shape == (m, 3) and out["prop_pos0"].shape[0] == m
    assert int(out["rounds"].max()) <= 3 and bool((out["rounds"] == out["valid"].sum(1)).all())
    assert bool((out["valid"][:, 1:] <= out["valid"][:, :-1]).all())                 # rounds are a prefix
    assert not bool((out["success"] & out["failed"]).any())
    assert int(out["success"].sum()) > 0.3 * m
    used = out["actions"][out["valid"]]
    assert bool((used[:, 0, 2] > 0.3).all()) and bool(torch.isfinite(used).all())   # pick heights on the table
    assert int((out["status"] & 7).ne(0).sum()) == 0
