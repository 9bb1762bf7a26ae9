"""Scripted expert of the colour-sorting task, batched over scenes.

Reference behaviour mirrored (mujoco_robot_environments/tasks/rearrangement.py):
  * ``props_info``              :227-295  prop geom world pose (``geom_xpos`` / ``geom_xmat``), rgba, colour label
  * ``prop_pick(prop_id)``      :579-595  grasp pose = prop position + mju_mat2Quat(R_xyz(0, 180 deg, yaw)), yaw from the prop's xmat
  * ``prop_place(prop_id, ..)`` :597-665  rejection-sample a position in [min_pose, max_pose] on a COPY of the physics until
                                          ``forward()`` lists no contact with dist <= 0.05 between the prop and anything but the table
  * ``sort_colours()``          :700-751  first prop (in prop order) outside the zone its colour maps to -> (True, pick, place)
  * camera maths                :480-548  ``_get_camera_intrinsics`` / ``_get_camera_extrinsics`` / ``world_2_pixel`` / ``pixel_2_world``
and environment/props.py:13-20 (COLOURS), :286-291 (colour noise), :604-617 (colour of prop i).

Everything here is small tensor algebra on the scene batch (torch = plumbing); the rejection loop behind ``prop_place`` runs on the
device in one launch (``b2p_place_search``: kinematics + the step kernel's broad/narrow phase per proposal).
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Sequence, Tuple

import numpy as np
import torch

# environment/props.py:13-20
COLOURS: Dict[str, Tuple[float, float, float, float]] = {
    "red": (1.0, 0.0, 0.0, 1.0), "green": (0.0, 1.0, 0.0, 1.0), "blue": (0.0, 0.0, 1.0, 1.0),
    "yellow": (1.0, 1.0, 0.0, 1.0), "cyan": (0.0, 1.0, 1.0, 1.0), "magenta": (1.0, 0.0, 1.0, 1.0),
}
COLOUR_NAMES = list(COLOURS)
DOWN_QUAT = (0.0, 0.0, 1.0, 0.0)     # mju_mat2Quat(R.from_euler('xyz', [0, 180, 0], degrees=True)), rearrangement.py:641-643


def sample_prop_colours(n_env: int, max_props: int, colours: Sequence[str], rng: np.random.Generator, colour_noise: float = 0.1):
    """props.py:604-617 + :286-291 per scene: prop 0 / 1 take ``colours[0]`` / ``colours[1]``, later props a uniformly random
    entry; rgba = COLOURS[colour] + U(-noise, noise) per channel, alpha 1.  -> (colour index into COLOUR_NAMES [n, P] int64,
    rgba [n, P, 4] float32).  The reference draws from the unseeded global ``random`` / ``np.random``; here ``rng`` is explicit."""
    for c in colours:
        if c not in COLOURS:
            raise ValueError("Unknown colour %r" % c)
    pick = rng.integers(0, len(colours), (n_env, max_props))
    pick[:, :2] = np.arange(min(2, max_props))[None, :] % len(colours)
    idx = np.array([COLOUR_NAMES.index(c) for c in colours])[pick]
    base = np.array([COLOURS[c] for c in COLOUR_NAMES])[idx]
    rgba = base + rng.uniform(-colour_noise, colour_noise, base.shape)
    rgba[..., 3] = 1.0
    return idx.astype(np.int64), rgba.astype(np.float32)


def zone_table(task_cfg: dict):
    """-> (lo [n_colour, 3], hi [n_colour, 3]) indexed by COLOUR_NAMES: the zone a colour maps to, with the reference's bounds
    ``location -+ size / 2`` in x, y and z fixed to 0.4 (``get_location_bounds``, rearrangement.py:703-710)."""
    lo, hi = np.zeros((len(COLOUR_NAMES), 3)), np.zeros((len(COLOUR_NAMES), 3))
    for i, c in enumerate(COLOUR_NAMES):
        t = task_cfg["target_locations"][task_cfg["colour_target_map"][c]]
        loc, size = t["location"], t["size"]
        lo[i] = (loc[0] - size[0] / 2, loc[1] - size[1] / 2, 0.4)
        hi[i] = (loc[0] + size[0] / 2, loc[1] + size[1] / 2, 0.4)
    return lo, hi


# ------------------------------------------------------------------------------------------ rotations (batched, torch)
def mat2quat(m: torch.Tensor) -> torch.Tensor:
    """mju_mat2Quat for [n, 9] row-major matrices -> [n, 4] wxyz (same branch order as the kernel / oracle)."""
    from .controllers import _mat2quat
    return _mat2quat(m)


def euler_xyz_matrix(rx: torch.Tensor, ry: torch.Tensor, rz: torch.Tensor) -> torch.Tensor:
    """scipy ``R.from_euler('xyz', [rx, ry, rz])`` (extrinsic x, then y, then z: R = Rz Ry Rx), radians -> [n, 9] row-major."""
    cx, sx, cy, sy, cz, sz = torch.cos(rx), torch.sin(rx), torch.cos(ry), torch.sin(ry), torch.cos(rz), torch.sin(rz)
    return torch.stack([cz * cy, cz * sy * sx - sz * cx, cz * sy * cx + sz * sx,
                        sz * cy, sz * sy * sx + cz * cx, sz * sy * cx - cz * sx,
                        -sy, cy * sx, cy * cx], -1)


def grasp_quat_from_xmat(xmat: torch.Tensor) -> torch.Tensor:
    """prop_pick (rearrangement.py:585-593): yaw = |deg(atan2(R10, R00))|; yaw = min(yaw, yaw - 90) (i.e. yaw - 90);
    grasp quat = mju_mat2Quat(R_xyz(0, 180, yaw) in degrees).  xmat [n, 9] -> quat [n, 4] wxyz."""
    yaw = torch.rad2deg(torch.atan2(xmat[:, 3], xmat[:, 0])).abs()
    yaw = torch.minimum(yaw, yaw - 90.0)
    z = torch.zeros_like(yaw)
    return mat2quat(euler_xyz_matrix(z, z + math.pi, torch.deg2rad(yaw)))


# ------------------------------------------------------------------------------------------ camera maths
class Camera:
    """The static overhead camera (transporter_data_collection.yaml:1-6) and the reference's pixel <-> world maps."""

    def __init__(self, cam_cfg: dict):
        from .compile.mjcf import quat_normalize, quat_to_mat
        self.name = cam_cfg["name"] + "/" + cam_cfg["name"]
        self.height, self.width, self.fovy = int(cam_cfg["height"]), int(cam_cfg["width"]), float(cam_cfg["fovy"])
        self.xpos = np.asarray(cam_cfg["pos"], np.float64)
        self.xmat = np.asarray(quat_to_mat(quat_normalize(np.asarray(cam_cfg["quat"], np.float64)))).reshape(3, 3)

    def intrinsics(self) -> np.ndarray:
        """_get_camera_intrinsics (rearrangement.py:480-492)."""
        f = (1.0 / np.tan(np.deg2rad(self.fovy) / 2)) * self.height / 2.0
        return np.array([[-f, 0, (self.width - 1) / 2], [0, f, (self.height - 1) / 2], [0, 0, 1]])

    def extrinsics(self) -> np.ndarray:
        """_get_camera_extrinsics (rearrangement.py:494-505): world -> camera."""
        e = np.eye(4)
        e[:3, :3] = self.xmat.T
        e[:3, 3] = -self.xmat.T @ self.xpos
        return e

    def world_2_pixel(self, coords) -> torch.Tensor:
        """rearrangement.py:537-550 for [n, 3] world points -> int32 [n, 2] (u, v)."""
        c = torch.as_tensor(coords)
        K = torch.as_tensor(self.intrinsics(), dtype=torch.float64, device=c.device)
        E = torch.as_tensor(self.extrinsics(), dtype=torch.float64, device=c.device)
        cam = c.to(torch.float64) @ E[:3, :3].T + E[:3, 3]
        img = cam @ K.T
        return torch.round(img[:, :2] / img[:, 2:3]).to(torch.int32)

    def pixel_2_world(self, coords, depth_val) -> torch.Tensor:
        """rearrangement.py:507-535 for [n, 2] pixels and the rendered depth at those pixels [n] -> [n, 3]."""
        c = torch.as_tensor(coords).to(torch.float64)
        d = torch.as_tensor(depth_val, device=c.device).to(torch.float64)
        Ki = torch.as_tensor(np.linalg.inv(self.intrinsics()), dtype=torch.float64, device=c.device)
        Ei = torch.as_tensor(np.linalg.inv(self.extrinsics()), dtype=torch.float64, device=c.device)
        cam = (torch.cat([c, torch.ones_like(c[:, :1])], 1) @ Ki.T) * (-d[:, None])
        return cam @ Ei[:3, :3].T + Ei[:3, 3]

    def metadata(self) -> dict:
        """get_camera_metadata (rearrangement.py:552-577); quaternion xyzw of extrinsics[:3,:3] like scipy's as_quat, and the
        reference's (transposed) indexing ``extrinsics[3, 0..2]`` for x, y, z."""
        K, E = self.intrinsics(), self.extrinsics()
        q = mat2quat(torch.as_tensor(E[:3, :3].reshape(1, 9)))[0].numpy()
        return {"intrinsics": {"fx": K[0, 0], "fy": K[1, 1], "cx": K[0, 2], "cy": K[1, 2]},
                "extrinsics": {"x": E[3, 0], "y": E[3, 1], "z": E[3, 2], "qx": q[1], "qy": q[2], "qz": q[3], "qw": q[0]}}


# ------------------------------------------------------------------------------------------ the expert
class SortColoursExpert:
    """``sort_colours`` / ``prop_pick`` / ``prop_place`` of a ``BatchedRearrangementEnv`` (kept outside the env class so
    that the env module stays the reset/step hot path)."""

    MAX_PLACE_ATTEMPTS = 10000         # rearrangement.py:629
    PLACE_DIST = 0.05                  # rearrangement.py:623

    def __init__(self, env):
        self.env = env
        task = env._cfg["task"]
        if "target_locations" not in task:
            raise ValueError("sort_colours needs task=rearrangement_w_targets (target_locations / colour_target_map)")
        lo, hi = zone_table(task)
        kw = dict(device=env.device, dtype=env.dtype)
        self.zone_lo, self.zone_hi = torch.tensor(lo, **kw), torch.tensor(hi, **kw)
        seed = int(task["initializers"]["seed"]) + 1           # prop_place_random_state, rearrangement.py:208
        self._rs = np.random.RandomState(seed=seed) if env.num_envs == 1 else None
        self._gen = torch.Generator(device=env.device); self._gen.manual_seed(seed)
        self._seed, self._place_calls = int(seed), 0
        self.place_attempts = 0
        self.place_failed = torch.zeros(env.num_envs, dtype=torch.bool, device=env.device)

    # -- props_info
    def prop_geom_pose(self):
        p = self.env.physics
        gid = self.env._prop_gid
        return p.data.geom_xpos[:, gid], p.data.geom_xmat[:, gid]

    def within_target(self) -> torch.Tensor:
        """is_within_target per scene / prop -> bool [n, P] (closed bounds, x and y only)."""
        pos, _ = self.prop_geom_pose()
        lo, hi = self.zone_lo[self.env.prop_colour], self.zone_hi[self.env.prop_colour]
        return ((pos[..., :2] >= lo[..., :2]) & (pos[..., :2] <= hi[..., :2])).all(-1)

    def prop_pick(self, prop_id: torch.Tensor) -> torch.Tensor:
        pos, mat = self.prop_geom_pose()
        e = torch.arange(self.env.num_envs, device=self.env.device)
        return torch.cat([pos[e, prop_id], grasp_quat_from_xmat(mat[e, prop_id])], 1)

    def _uniform(self, lo: torch.Tensor, hi: torch.Tensor) -> torch.Tensor:
        """``distributions.Uniform(min_pose, max_pose)(random_state)``: one (x, y, z) draw per scene and attempt.  A single
        scene draws from ``np.random.RandomState(seed + 1)`` exactly like the reference (rearrangement.py:208,638); a batch
        draws on the device (no host round trip inside the rejection loop)."""
        if self._rs is not None:
            return torch.tensor(self._rs.uniform(lo[0].cpu().numpy().astype(np.float64), hi[0].cpu().numpy().astype(np.float64))[None], device=lo.device, dtype=lo.dtype)
        return lo + (hi - lo) * torch.rand(lo.shape, device=lo.device, dtype=lo.dtype, generator=self._gen)

    def prop_place(self, prop_id: torch.Tensor, min_pose: torch.Tensor, max_pose: torch.Tensor, active: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Collision-free place pose per scene -> [n, 7]; scenes outside ``active`` get zeros.  The search runs in the kernel's
        own scratch copy of the scene (``deepcopy(self._physics)`` in the reference): state and derived outputs are not written.
        A scene that finds no pose within MAX_PLACE_ATTEMPTS raises for a strict (single-scene) env like the reference
        (rearrangement.py:665) and is flagged in ``self.place_failed`` (and ``env.failed``) for a batch."""
        env, p = self.env, self.env.physics
        n = env.num_envs
        todo = torch.ones(n, dtype=torch.bool, device=env.device) if active is None else active.to(env.device, torch.bool)
        quat = torch.tensor(DOWN_QUAT, device=env.device, dtype=env.dtype).expand(n, 4)
        self._place_calls += 1
        # one launch: every scene runs its own rejection loop on the device (proposal -> kinematics -> full collision pass ->
        # the dist <= PLACE_DIST rule of :623), the bound state is never written
        res = p.place_search(prop_id, min_pose, max_pose, env._table_gid, self.PLACE_DIST, self.MAX_PLACE_ATTEMPTS,
                             seed=self._seed * 1000003 + self._place_calls, mask=todo)
        found = todo & (res[:, 3] > 0)
        self.place_attempts += int(res[:, 3].abs().max().item()) if bool(todo.any()) else 0
        out = torch.cat([res[:, :3], quat], 1) * found[:, None]
        todo = todo & ~found
        self.place_failed = todo
        if bool(todo.any()):
            if env.strict:
                raise RuntimeError("Failed to find collision free place pose.")          # rearrangement.py:665
            env.failed |= todo
        return out

    def sort_colours(self):
        """-> (in_progress bool [n], pick_pose [n, 7], place_pose [n, 7], prop_id int64 [n]); rows of finished scenes are zero."""
        env = self.env
        cand = env.has_prop & ~self.within_target()
        in_progress = cand.any(1)
        prop_id = torch.argmax(cand.to(torch.int8), dim=1)             # first misplaced prop in prop order
        pick = self.prop_pick(prop_id) * in_progress[:, None]
        colour = env.prop_colour[torch.arange(env.num_envs, device=env.device), prop_id]
        place = self.prop_place(prop_id, self.zone_lo[colour], self.zone_hi[colour], in_progress)
        in_progress = in_progress & ~self.place_failed
        return in_progress, pick * in_progress[:, None], place, prop_id
