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


# ---------------------------------------------------------------------------------------------- the drawing scene's render model
def _draw_assets():
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "mujoco_robot_environments_b200", "assets")
    with open(os.path.join(root, "lasa_draw.b2pm"), "rb") as f:
        pb = f.read()
    with open(os.path.join(root, "lasa_draw.b2pr"), "rb") as f:
        rb = f.read()
    with open(os.path.join(root, "lasa_draw.names.json")) as f:
        names = json.load(f)
    return pb, rb, names


def test_draw_render_blob_is_reproducible_and_the_tool_piece_is_a_tight_prism():
    pb, rb, names = _draw_assets()
    assert R.build_render_blob(pb, names["geom"], C.LASA_CAMERAS[0], C.LASA_ZONES, colours=C.LASA_COLOURS) == rb
    assert rnd.header(rb) == dict(ngeom=13, nplane=rnd.header(rb)["nplane"], nzone=1, W=640, H=640, nprop=0)
    pl = R.hull_planes(R.cylinder_prism(0.015, 0.05), max_planes=None)
    assert len(pl) == R.CYLINDER_SIDES + 2
    ang = np.linspace(0, 2 * np.pi, 1000)
    surf = np.concatenate([np.stack([0.015 * np.cos(ang), 0.015 * np.sin(ang), np.full_like(ang, z)], 1) for z in (-0.05, 0.0, 0.05)], 0)
    slackness = (surf @ pl[:, :3].T - pl[:, 3]).max(1)           # how far outside the polytope a cylinder point lies (<= 0: inside)
    assert slackness.max() <= 1e-12 and slackness.min() > -2.5e-5  # circumscribed: the cylinder touches every side face, sticks in by < 0.02 mm


def test_draw_scene_known_answers():
    """main_camera (config/arena/cameras/lasa.yaml) at the arm's home pose: a table pixel's depth is the ray / plane z = 0.4 answer of
    the pinhole model; the tool piece, the links and the draw plane's zone are in view."""
    pb, rb, names = _draw_assets()
    o = Oracle(pb); o.reset(); o.qpos[:7] = [0, -0.785, 0, -2.356, 0, 1.571, 0.785]; o.forward()
    rgb, depth, seg = rnd.render(rb, np.array(o.geom_xpos), np.array(o.geom_xmat), None, None, None)
    tool, table = names["geom"].index("panda nohand/tool"), names["geom"].index("table/table")
    assert (seg == tool).sum() > 50 and (seg == table).sum() > 100000 and (seg == 13).sum() > 10000      # 13 = ngeom + zone 0
    cam = C.LASA_CAMERAS[0]
    q = np.asarray(cam["quat"], float); q /= np.linalg.norm(q)
    w, x, y, z = q
    M = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)], [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                  [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
    f = (1 / np.tan(np.deg2rad(cam["fovy"]) / 2)) * cam["height"] / 2
    vs, us = np.nonzero(seg == table)
    for k in (0, len(vs) // 2, len(vs) - 1):
        u, v = us[k], vs[k]
        d = M[:, 0] * ((u - (cam["width"] - 1) / 2) / f) + M[:, 1] * (((cam["height"] - 1) / 2 - v) / f) - M[:, 2]
        t = (0.4 - cam["pos"][2]) / d[2]                                                               # axial depth: the direction has -1 along the view axis
        assert abs(depth[v, u] - t) < 2e-5 * t, (u, v, depth[v, u], t)
    assert tuple(rgb[vs[0], us[0]]) != tuple(rgb[(seg == tool)][0])                                    # white table, petrol tool piece
