"""Render-model blob ("B2PR") for the CUDA ray-cast observations.

Derived from the physics blob (geom types / sizes / convex hulls / static poses), the scene config (overhead camera of
config/arena/cameras/transporter_data_collection.yaml:1-6, target zones of task/rearrangement_w_targets.yaml:12-56 added as
visual boxes at tasks/rearrangement.py:104-116) and the lights of models/arenas/empty_assets/arena.xml:14-16.  The reference
renders the arm's group-2 visual meshes through OpenGL (unavailable on this box); the ray-caster renders the collision hulls
of the same links, coloured with the material of the link's visual geoms (panda_nohand.xml:26-30, 2f85.xml:7-10,46-53).

Layout (little endian): int32 header[16] | float32 cam[32] | per geom GEOM_WORDS x 4 bytes | zones ZONE_WORDS x float32 | planes float32[4].
``include/b2p_render.h`` documents the same layout for C.
"""
from __future__ import annotations

from typing import Optional, Sequence

import numpy as np

from . import blob as B
from .mjcf import quat_normalize, quat_to_mat

MAGIC = 0x52503242      # "B2PR"
VERSION = 1
GEOM_WORDS = 32         # per geom: 8 int32 + 24 float32
ZONE_WORDS = 12         # pos3 half3 rgba4 pad2
AABB_PAD = 1e-4         # hull AABB pre-test is inflated so that it can never reject a ray the planes accept

# material colour of the link a collision geom belongs to (first match wins)
_COLOURS = [
    ("right_pad1", (0.55, 0.55, 0.55)), ("left_pad1", (0.55, 0.55, 0.55)), ("right_pad2", (0.45, 0.45, 0.45)), ("left_pad2", (0.45, 0.45, 0.45)),
    ("driver", (0.4627, 0.4627, 0.4627)), ("robotiq_2f85", (0.149, 0.149, 0.149)), ("panda nohand/link", (1.0, 1.0, 1.0)),
    ("table/table", (0.5, 0.5, 0.5)), ("ground", (0.2, 0.3, 0.4)), ("prop_", (0.5, 0.5, 0.5)),
]
GROUND_RGB2 = (0.1, 0.2, 0.3)       # arena.xml:5 checker rgb2; texrepeat 5 over a 2x2 builtin checker => 0.1 m squares
HEADLIGHT = dict(ambient=0.1, diffuse=0.4)                      # MuJoCo visual/headlight defaults (specular terms are not modelled)
SCENE_LIGHT = dict(pos=(0.7, 0.0, 1.6), diffuse=0.7)            # arena.xml:16 (positional, castshadow=false)
ZFAR = 100.0


MAX_HULL_PLANES = 128   # hulls with more facets than this are rendered as the discrete oriented polytope below ...
SMALL_HULL_PLANES = 64  # ... and small parts (bounding radius <= SMALL_HULL_RADIUS: the gripper linkage) with fewer directions:
SMALL_HULL_RADIUS = 0.045  # either way the polytope sticks out of the exact hull by less than ~1 mm (under half a pixel)


def _kdop_directions(n: int) -> np.ndarray:
    """The 6 axis directions (so the polytope stays inside the hull's AABB, which the kernel uses as a pre-test) plus a
    Fibonacci lattice on the sphere."""
    m = n - 6
    i = np.arange(m) + 0.5
    z = 1 - 2 * i / m
    phi = np.pi * (1 + 5 ** 0.5) * i
    fib = np.stack([np.sqrt(1 - z * z) * np.cos(phi), np.sqrt(1 - z * z) * np.sin(phi), z], 1)
    return np.concatenate([np.eye(3), -np.eye(3), fib], 0)


def _spread(n: int) -> np.ndarray:
    """Fixed permutation that visits 0..n-1 in a low-discrepancy order, so that the first few planes of a hull already face
    every direction and a ray that misses the hull is rejected after a handful of planes."""
    return np.argsort((np.arange(n) * 0.6180339887498949) % 1.0, kind="stable")


def hull_planes(verts: np.ndarray, max_planes="auto") -> np.ndarray:
    """Supporting planes (n, d) with n.x <= d inside.  Exact facets (coplanar triangles merged) when there are at most
    ``max_planes`` of them; otherwise the tightest polytope with ``max_planes`` fixed face directions around the hull (outer
    approximation; the deviation stays under ~1 mm, less than half a pixel from the camera).  Planes come out in a spread order."""
    from scipy.spatial import ConvexHull
    eq = ConvexHull(verts).equations          # n.x + off <= 0 inside
    key = np.round(eq / 1e-7).astype(np.int64)
    _, idx = np.unique(key, axis=0, return_index=True)
    eq = eq[np.sort(idx)]
    if max_planes == "auto":
        c = (verts.min(0) + verts.max(0)) / 2
        max_planes = SMALL_HULL_PLANES if np.linalg.norm(verts - c, axis=1).max() <= SMALL_HULL_RADIUS else MAX_HULL_PLANES
    if max_planes is None or len(eq) <= max_planes:
        pl = np.concatenate([eq[:, :3], -eq[:, 3:4]], 1)
    else:
        dirs = _kdop_directions(max_planes)
        pl = np.concatenate([dirs, (verts @ dirs.T).max(0)[:, None]], 1)
    return pl[_spread(len(pl))]


CYLINDER_SIDES = 64     # a cylinder geom is rendered as the circumscribed 64-gon prism (66 planes; sticks out by r (1 / cos(pi/64) - 1) = 0.02 mm at r = 15 mm)


def cylinder_prism(radius: float, half_height: float, sides: int = CYLINDER_SIDES) -> np.ndarray:
    """Vertices of the prism whose side faces are tangent to the cylinder (axis z)."""
    ang = 2 * np.pi * (np.arange(sides) + 0.5) / sides
    rc = radius / np.cos(np.pi / sides)
    ring = np.stack([rc * np.cos(ang), rc * np.sin(ang)], 1)
    return np.concatenate([np.concatenate([ring, np.full((sides, 1), z)], 1) for z in (-half_height, half_height)], 0)


def build_render_blob(model_blob: bytes, geom_names: Sequence[str], camera: dict, zones: Optional[dict] = None,
                      colours: Optional[Sequence] = None) -> bytes:
    """``colours``: (name fragment, rgb) pairs looked up before the rearrangement scene's table of materials."""
    counts, _, A = B.unpack(model_blob)
    palette = list(colours or []) + _COLOURS
    ng = counts["ngeom"]
    geoms = np.zeros((ng, GEOM_WORDS), np.float32)
    gi = geoms.view(np.int32)
    planes, mesh_planes = [], {}
    prop_of = {int(g): p for p, g in enumerate(A["prop_geomid"])}
    for g in range(ng):
        t, mid = int(A["geom_type"][g]), int(A["geom_meshid"][g])
        adr = num = 0
        if t == B.GEOM_MESH:
            if mid not in mesh_planes:
                a, n = int(A["mesh_vertadr"][mid]), int(A["mesh_vertnum"][mid])
                pl = hull_planes(np.asarray(A["mesh_vert"]).reshape(-1, 3)[a:a + n])
                mesh_planes[mid] = (sum(len(p) for p in planes), len(pl))
                planes.append(pl)
            adr, num = mesh_planes[mid]
        aabb = np.asarray(A["geom_aabb"][g], np.float64)
        size = np.asarray(A["geom_size"][g], np.float64)
        if t == B.GEOM_CYLINDER:      # no cylinder primitive in the ray-caster: a convex polytope like the link hulls
            pl = hull_planes(cylinder_prism(size[0], size[1]), max_planes=None)
            adr, num = sum(len(p) for p in planes), len(pl)
            planes.append(pl)
            t = B.GEOM_MESH
            aabb = np.array([0, 0, 0, size[0] / np.cos(np.pi / CYLINDER_SIDES), size[0] / np.cos(np.pi / CYLINDER_SIDES), size[1]])
        static = int(A["geom_bodyid"][g]) == 0
        gi[g, 0:8] = (t, adr, num, prop_of.get(g, -1), 1 if static else 0, 0, 0, 0)
        rgb = next(c for k, c in palette if k in geom_names[g])
        f = geoms[g]
        f[8:11] = rgb; f[11] = 1.0
        f[12:15] = size if t != B.GEOM_MESH else aabb[3:6] + AABB_PAD
        f[15:18] = aabb[0:3] if t == B.GEOM_MESH else 0.0
        f[18] = float(A["geom_rbound"][g]) if t != B.GEOM_PLANE else 0.0
        if static:
            f[19:22] = np.asarray(A["geom_pos"][g]); f[22:31] = quat_to_mat(quat_normalize(np.asarray(A["geom_quat"][g]))).reshape(9)
    zl = []
    for z in (zones or {}).values():
        zl.append(list(z["location"]) + list(z["size"]) + list(z["rgba"]) + [0.0, 0.0])
    zarr = np.asarray(zl, np.float32).reshape(-1, ZONE_WORDS)
    parr = np.concatenate(planes, 0).astype(np.float32) if planes else np.zeros((0, 4), np.float32)
    cam = np.zeros(32, np.float32)
    h, w = int(camera["height"]), int(camera["width"])
    cam[0:3] = camera["pos"]
    cam[3:12] = quat_to_mat(quat_normalize(np.asarray(camera["quat"], np.float64))).reshape(9)
    cam[12] = (1.0 / np.tan(np.deg2rad(float(camera["fovy"])) / 2)) * h / 2.0      # _get_camera_intrinsics, rearrangement.py:485
    cam[13], cam[14], cam[15] = (w - 1) / 2, (h - 1) / 2, ZFAR
    cam[16], cam[17] = HEADLIGHT["ambient"], HEADLIGHT["diffuse"]
    cam[18:21] = SCENE_LIGHT["pos"]; cam[21] = SCENE_LIGHT["diffuse"]
    cam[22:25] = GROUND_RGB2; cam[25] = 0.1
    header = np.zeros(16, np.int32)
    header[:8] = (MAGIC, VERSION, ng, len(parr), len(zarr), w, h, counts["nprop"])
    return header.tobytes() + cam.tobytes() + geoms.tobytes() + zarr.tobytes() + parr.tobytes()
