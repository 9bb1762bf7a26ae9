"""CPU tests of the render model and of the plain-C restatement of the overhead-camera ray-cast.  The reference's
GL images are unavailable on this box, so the oracle is pinned by analytic known answers: pinhole geometry of
`_get_camera_intrinsics/_extrinsics` (rearrangement.py:480-505), depth of the table / zone / prop tops, hand-computed shading."""
import json
import os

import numpy as np
import pytest

import rnd
import scenarios as sc
from orc import Oracle
from mujoco_robot_environments_b200 import config as C
from mujoco_robot_environments_b200.compile import blob as B
from mujoco_robot_environments_b200.compile import render as R

ASSETS = os.path.join(sc.ROOT, "mujoco_robot_environments_b200", "assets")


REACH = [0.0, 0.35, 0.0, -1.9, 0.0, 2.25, 0.785]      # arm stretched over the table (at home it is outside the overhead view, rearrangement.py:327)


def _scene(seed=5, arm=None):
    st, size, active = sc.make_states(1, seed=seed)
    if arm is not None:
        st[0, :7] = arm
    o = Oracle(sc.default_blob()); sc.oracle_load(o, st[0], size[0], active[0]); o.forward()
    return o, st[0], size[0], active[0]


def test_render_blob_is_reproducible_from_the_physics_blob():
    names = json.load(open(os.path.join(ASSETS, "rearrangement_n4.names.json")))
    rb = R.build_render_blob(sc.default_blob(), names["geom"], C.CAMERAS[0], C.TASK_GROUPS["rearrangement_w_targets"]["target_locations"])
    assert rb == rnd.render_blob()
    h = rnd.header(rb)
    assert (h["W"], h["H"], h["nzone"], h["ngeom"]) == (640, 480, 6, 30)


def test_hull_planes_support_their_vertices():
    counts, _, A = B.unpack(sc.default_blob())
    v = np.asarray(A["mesh_vert"]).reshape(-1, 3)
    for mid in range(counts["nmesh"]):
        a, n = int(A["mesh_vertadr"][mid]), int(A["mesh_vertnum"][mid])
        exact = R.hull_planes(v[a:a + n], max_planes=None)
        s = v[a:a + n] @ exact[:, :3].T - exact[:, 3]
        assert s.max() < 1e-9                                   # every vertex inside every half space
        assert ((np.abs(s) < 1e-7).sum(0) >= 3).all()           # every plane is a facet
        pl = R.hull_planes(v[a:a + n])
        assert len(pl) <= max(R.MAX_HULL_PLANES, len(exact) if len(exact) <= R.MAX_HULL_PLANES else 0) and np.abs(np.linalg.norm(pl[:, :3], axis=1) - 1).max() < 1e-9
        s = v[a:a + n] @ pl[:, :3].T - pl[:, 3]
        assert s.max() < 1e-9 and (np.abs(s).min(0) < 1e-9).all()      # outer approximation, every plane touches the hull
        if len(exact) > R.MAX_HULL_PLANES:
            # how far the polytope sticks out of the exact hull: its vertices are bounded by the support gap in the exact facet directions
            lo, hi = v[a:a + n].min(0), v[a:a + n].max(0)
            for k in range(3):                                  # the axis planes are among the directions: polytope inside the AABB
                assert np.abs(pl[np.argmax(pl[:, k]), 3] - hi[k]) < 1e-12 and np.abs(pl[np.argmin(pl[:, k]), 3] + lo[k]) < 1e-12


def test_oracle_known_answers():
    o, st, size, active = _scene()
    blob = rnd.render_blob(); h = rnd.header(blob)
    cam = __import__("mujoco_robot_environments_b200.expert", fromlist=["Camera"]).Camera(C.CAMERAS[0])
    rgba = np.tile(np.array([[0.9, 0.1, 0.2, 1.0]], np.float32), (4, 1))
    import torch
    # a bare table pixel: depth = camera height - table top, seg = table geom
    u, v = cam.world_2_pixel(torch.tensor([[0.8, -0.55, 0.4]]))[0].tolist()
    win = (u - 2, v - 2, u + 3, v + 3)
    rgb, depth, seg = rnd.render(blob, o.geom_xpos, o.geom_xmat, rgba, size, active, window=win)
    assert seg[v, u] == 1 and abs(depth[v, u] - 0.9) < 1e-6
    # hand-computed Lambert shading of that pixel: headlight (0.1 + 0.4 n.back) + point light 0.7 n.l
    K = cam.intrinsics(); f = K[1, 1]
    d = cam.xmat @ np.array([(u - K[0, 2]) / f, (K[1, 2] - v) / f, -1.0]); p = cam.xpos + 0.9 * d
    L = np.array([0.7, 0.0, 1.6]) - p; s = 0.1 + 0.4 * 1.0 + 0.7 * (L[2] / np.linalg.norm(L))
    assert abs(int(rgb[v, u, 0]) - int(min(1.0, 0.5 * s) * 255 + 0.5)) <= 1 and rgb[v, u, 0] == rgb[v, u, 1] == rgb[v, u, 2]
    # centre of a target zone: translucent layer on top (depth 1.3 - 0.41), seg = ngeom + zone index, colour tinted towards the zone's
    zones = list(C.TASK_GROUPS["rearrangement_w_targets"]["target_locations"].values())
    zc = np.array(zones[4]["location"])            # top_middle (magenta), away from the arm
    u, v = cam.world_2_pixel(torch.tensor(zc[None]))[0].tolist()
    rgb, depth, seg = rnd.render(blob, o.geom_xpos, o.geom_xmat, rgba, size, np.zeros(4, np.uint8), window=(u, v, u + 1, v + 1))
    assert seg[v, u] == h["ngeom"] + 4 and abs(depth[v, u] - (1.3 - 0.41)) < 1e-6
    assert rgb[v, u, 0] > rgb[v, u, 1] and rgb[v, u, 2] > rgb[v, u, 1]
    # every active prop: its projected centre shows the prop (seg id), at the depth of its top face, with the per-scene colour
    for k in range(4):
        if not active[k]:
            continue
        g = 26 + k
        c = o.geom_xpos[g]
        u, v = cam.world_2_pixel(torch.tensor([[c[0], c[1], c[2] + size[k, 2]]]))[0].tolist()
        rgb, depth, seg = rnd.render(blob, o.geom_xpos, o.geom_xmat, rgba, size, active, window=(u, v, u + 1, v + 1))
        assert seg[v, u] == g
        assert abs(depth[v, u] - (1.3 - (c[2] + size[k, 2]))) < 1e-5
        assert rgb[v, u, 0] > 150 and rgb[v, u, 1] < 60
        # pixel_2_world through the rendered depth lands on the prop's top face (rearrangement.py:507-535)
        w = cam.pixel_2_world(torch.tensor([[float(u), float(v)]]), torch.tensor([float(depth[v, u])]))[0].numpy()
        assert abs(w[2] - (c[2] + size[k, 2])) < 1e-5 and np.abs(w[:2] - c[:2]).max() < 2.5e-3      # within a pixel (2.2 mm at 0.87 m)
    # at home the arm is outside the view; stretched over the table its link / gripper hulls are drawn, nearer than the table
    rgb, depth, seg = rnd.render(blob, o.geom_xpos, o.geom_xmat, rgba, size, active, window=(0, 256, 640, 480))
    assert not ((seg[256:] >= 2) & (seg[256:] <= 25)).any() and (seg[256:] >= 0).all()
    o2, _, _, _ = _scene(arm=REACH)
    rgb, depth, seg = rnd.render(blob, o2.geom_xpos, o2.geom_xmat, rgba, size, active, window=(0, 256, 640, 480))
    arm = (seg[256:] >= 2) & (seg[256:] <= 25)
    assert arm.sum() > 10000 and len(np.unique(seg[256:][arm])) >= 6
    assert depth[256:][arm].max() < 0.9 and depth[256:][arm].min() > 0.2
    assert rgb[256:][seg[256:] == 10].mean() > 150            # link6 hull: white panda material, lit from above
