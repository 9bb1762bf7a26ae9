"""BatchedRearrangementEnv: the reference's ``RearrangementEnv`` reset/step API with a leading scene dimension.

Reference behaviour mirrored (mujoco_robot_environments/tasks/rearrangement.py):
  * constructor  :54-210   scene = arena + table + FER/Robotiq2F85 + N props (N, sizes sampled per scene, props.py:277-291,604-617)
  * ``reset()``  :297-337  physics.reset, arm -> home, PropPlacer (environment/prop_initializer.py:164-283), RobotArm, eef_home_pose
  * ``step()``   :339-356  alternates scripted ``pick`` (:358-401) and ``place`` (:403-440): 5 x (set_target, run_controller)
The env batch is heterogeneous: every scene is padded to ``model.nprop`` props; props a scene does not
have are parked away from the table with collisions and dynamics disabled.

Camera observations of the reference come from ``mujoco.Renderer`` (OpenGL): unavailable on this box.  ``render()`` ray-casts
the overhead camera on the GPU instead (csrc/b2p_render.cu); ``_compute_observation`` returns the state observation (poses)
that the scripted expert consumes, so that stepping never pays for frames nobody reads.
"""
from __future__ import annotations

import collections
import enum
import math
from typing import Any, Dict, Iterable, Optional

import numpy as np
import torch

from . import config as cfglib
from . import ffi
from .controllers import BatchedMinMax, BatchedOSC
from .expert import Camera, SortColoursExpert, euler_xyz_matrix, mat2quat, sample_prop_colours
from .physics import BatchedPhysics
from .robot_arm import BatchedRobotArm

TABLE_GEOM = "table/table"
DOWN_QUAT = (0.0, 0.0, 1.0, 0.0)        # mju_mat2Quat(R.from_euler('xyz', [0, 180, 0])), rearrangement.py:391-393
PARK_XY = (5.0, 5.0)


class StepType(enum.IntEnum):           # dm_env.StepType
    FIRST = 0
    MID = 1
    LAST = 2


TimeStep = collections.namedtuple("TimeStep", ["step_type", "reward", "discount", "observation"])   # dm_env.TimeStep


class BatchedRearrangementEnv:
    def __init__(self, viewer=None, cfg: Optional[Dict[str, Any]] = None, num_envs: int = 1, device=None,
                 dtype: torch.dtype = torch.float32, overrides: Optional[Iterable[str]] = None, seed: Optional[int] = None,
                 sampler: str = "auto", strict: Optional[bool] = None, camera_obs: Optional[bool] = None):
        if viewer:
            raise RuntimeError("passive viewer: unavailable on this box (no graphics stack)")
        self._cfg = cfg if cfg is not None else cfglib.compose(overrides)
        self.num_envs = int(num_envs)
        self._physics = BatchedPhysics(self.num_envs, dtype=dtype, device=device)
        p = self._physics
        # what this class reads after a stepping launch: the end-effector site (observation) and the geom poses (expert, ray-caster);
        # contacts are only read after an explicit forward(), as in the reference (prop_initializer.py:221)
        p.set_out_fields("SITE", "GEOM")
        if abs(p.model.opt.timestep - float(self._cfg["physics_dt"])) > 1e-12:
            raise ValueError("physics_dt=%g differs from the compiled model's timestep %g: recompile the scene blob (tools/compile_scene.py)"
                             % (self._cfg["physics_dt"], p.model.opt.timestep))
        self.strict = (self.num_envs == 1) if strict is None else bool(strict)
        # the reference's observation IS the camera pair (rearrangement.py:460-478): on by default for the single-scene facade,
        # opt-in for a batch (2.15 MB per scene and observation)
        self.camera_obs = (self.num_envs == 1) if camera_obs is None else bool(camera_obs)
        self.device, self.dtype = p.device, p.dtype
        self.max_props = p.model.nprop
        pc = self._cfg["arena"]["props"]
        tseed = int(self._cfg["task"]["initializers"]["seed"]) if seed is None else int(seed)
        # scene construction randomness (props.py:604-607 num_objects, :277-284 cube size, :286-291 colour): unseeded in the reference
        self._crng = np.random.default_rng(tseed + 7919)
        self.n_props = torch.zeros(self.num_envs, dtype=torch.int64, device=self.device)
        self.has_prop = torch.zeros(self.num_envs, self.max_props, dtype=torch.bool, device=self.device)
        self.prop_half_size = torch.zeros(self.num_envs, self.max_props, 3, device=self.device, dtype=self.dtype)
        self.prop_type = torch.full((self.num_envs, self.max_props), ffi.GEOM["box"], dtype=torch.uint8, device=self.device)   # collider class per prop
        self.props_clamped = 0
        self.prop_colour = torch.zeros(self.num_envs, self.max_props, dtype=torch.int64, device=self.device)
        self.prop_rgba = torch.zeros(self.num_envs, self.max_props, 4, dtype=torch.float32, device=self.device)
        self.resample_scenes()
        self.camera = Camera(next(c for c in self._cfg["arena"]["cameras"] if c["name"] == "overhead_camera"))
        self.overhead_camera_height, self.overhead_camera_width = self.camera.height, self.camera.width
        self._expert: Optional[SortColoursExpert] = None
        self._renderer = None
        # controllers (hydra partial `controller(physics, arm)` in the reference, robot_arm.py:37,46)
        cc = self._cfg["robots"]["arm"]["controller_config"]["controller_params"]
        ee = self._cfg["robots"]["end_effector"]["controller_config"]
        self._osc = BatchedOSC(p, None, cc)
        self._minmax = BatchedMinMax(ee["max_val"], ee["min_val"], self.num_envs, self.device)
        self._robot: Optional[BatchedRobotArm] = None
        # PropPlacer randomness: np.random.RandomState(seed) per scene (rearrangement.py:207); scene e uses seed + e so that
        # scene 0 reproduces the reference's draw order exactly
        # "device": proposals drawn with the device generator, no host round trip inside the rejection loops
        self._sampler = sampler if sampler != "auto" else ("per_env" if self.num_envs <= 4096 else "device")
        self._dgen = torch.Generator(device=self.device); self._dgen.manual_seed(tseed)
        self._rs = [np.random.RandomState(seed=tseed + e) for e in range(self.num_envs)] if self._sampler == "per_env" else None
        self._vrng = np.random.default_rng(tseed)
        self._place_seed = (tseed * 2654435761) & 0x7FFFFFFF   # counter-based proposals of the device-side placer: a new stream per launch
        self._table_gid = p.geom_id(TABLE_GEOM)
        self._prop_gid = torch.tensor(p.model.prop_geomid, device=self.device)
        self.mode = None
        self.failed = torch.zeros(self.num_envs, dtype=torch.bool, device=self.device)
        self._active = None                                   # scenes being stepped by the current phase table (None = all)
        self.eef_home_pose = None

    def resample_scenes(self, mask: Optional[torch.Tensor] = None):
        """Draw prop count, cube size and colour / rgba for the scenes in ``mask`` (all by default): what constructing a NEW
        ``RearrangementEnv`` does in the reference (``add_objects``, props.py:583-639; the data generator builds a fresh env every
        10 episodes, transporter_network_data_generation.py:99-101).  Takes effect at the next ``reset`` of those scenes."""
        pc = self._cfg["arena"]["props"]
        idx = np.arange(self.num_envs) if mask is None else torch.nonzero(mask.to(self.device)).flatten().cpu().numpy()
        k = len(idx)
        if k == 0:
            return
        lo, hi = int(pc["min_objects"]), int(pc["max_objects"])
        n = np.full(k, lo) if lo == hi else self._crng.integers(lo, hi, k)                    # np.random.randint(lo, hi): hi exclusive
        if lo > self.max_props:
            raise ValueError("prop set needs at least %d props, the compiled model has %d" % (lo, self.max_props))
        # itl_rearrangement draws 4 or 5 props; the compiled model class holds 4 (MAXP): a fifth is not placed, and counted here
        self.props_clamped += int((n > self.max_props).sum())
        n = np.minimum(n, self.max_props)
        # shape per prop (add_objects, props.py:604-617: random.choice(shapes)) -> collider class and its size parameters
        #   cube: 3 x [U]; rectangle: U(3) (props.py:277-284); cylinder: (radius, half height) = U(2) (:360-364); apple: the asset's sphere
        shapes = list(pc.get("shapes", ["cube"]))
        unknown = [x for x in shapes if x not in cfglib.SHAPES]
        if unknown:
            raise ValueError("prop shapes %r are not among the collider classes %r" % (unknown, cfglib.SHAPES))
        sidx = self._crng.integers(0, len(shapes), (k, self.max_props))
        u = self._crng.uniform(pc["min_object_size"], pc["max_object_size"], (k, self.max_props, 3))
        half = np.zeros((k, self.max_props, 3)); ptype = np.full((k, self.max_props), ffi.GEOM["box"], np.uint8)
        for si, name in enumerate(shapes):
            sel = sidx == si
            if name == "cube":
                half[sel] = u[sel][:, :1].repeat(3, axis=1)
            elif name == "rectangle":
                half[sel] = u[sel]
            elif name == "cylinder":
                half[sel] = np.concatenate([u[sel][:, :2], np.zeros((int(sel.sum()), 1))], 1); ptype[sel] = ffi.GEOM["cylinder"]
            else:
                half[sel] = [cfglib.APPLE_RADIUS, 0.0, 0.0]; ptype[sel] = ffi.GEOM["sphere"]
        # "grey" of single_block has no COLOURS entry in the reference either
        cidx, rgba = sample_prop_colours(k, self.max_props, [c for c in pc["colours"] if c != "grey"], self._crng)
        di = torch.as_tensor(idx, device=self.device)
        self.n_props[di] = torch.tensor(n, device=self.device)
        self.has_prop[di] = torch.arange(self.max_props, device=self.device)[None, :] < self.n_props[di][:, None]
        self.prop_half_size[di] = torch.tensor(half, device=self.device, dtype=self.dtype)
        self.prop_type[di] = torch.tensor(ptype, device=self.device)
        self.prop_colour[di] = torch.tensor(cidx, device=self.device)
        self.prop_rgba[di] = torch.tensor(rgba, device=self.device)
        self._physics.set_env_params(self.prop_half_size, self.has_prop.to(torch.uint8), self.prop_type)

    # ------------------------------------------------------------------ accessors
    @property
    def physics(self) -> BatchedPhysics:
        return self._physics

    @property
    def robot(self) -> BatchedRobotArm:
        return self._robot

    def close(self):
        self._physics.close()

    # ------------------------------------------------------------------ placement (PropPlacer.__call__)
    def _sample_poses(self, envs: np.ndarray):
        """position ~ U(min_pose, max_pose) then yaw = pi * U(0, 1) about z (rearrangement.py:189-197): 3 + 1 draws per try."""
        w = self._cfg["task"]["initializers"]["workspace"]
        lo, hi = np.asarray(w["min_pose"], float), np.asarray(w["max_pose"], float)
        if self._sampler == "per_env":
            pos = np.stack([self._rs[e].uniform(lo, hi) for e in envs]) if len(envs) else np.zeros((0, 3))
            ang = np.array([np.pi * self._rs[e].uniform(0, 1) for e in envs])
        else:
            pos = self._vrng.uniform(lo, hi, (len(envs), 3)); ang = np.pi * self._vrng.uniform(0, 1, len(envs))
        quat = np.stack([np.cos(ang / 2), np.zeros_like(ang), np.zeros_like(ang), np.sin(ang / 2)], 1) if len(envs) else np.zeros((0, 4))
        return pos, quat

    def _collides(self, k: int) -> torch.Tensor:
        """_has_collisions_with_prop (prop_initializer.py:121-140): any listed contact (dist <= 0.6, i.e. inside the 0.15
        detection margin) that involves prop k's geom and not the table -> bool [n_env]."""
        p = self._physics
        c = p.data.contact
        valid = torch.arange(p.maxcon, device=self.device)[None, :] < p.data.ncon[:, None]
        gk = int(self._prop_gid[k])
        touches = (c["geom1"] == gk) | (c["geom2"] == gk)
        table = (c["geom1"] == self._table_gid) | (c["geom2"] == self._table_gid)
        return (valid & touches & ~table & (c["dist"] <= 0.6)).any(dim=1)

    PLACEMENT_RESTARTS = 8

    def _place_props(self, mask: torch.Tensor):
        """PropPlacer.__call__ (prop_initializer.py:164-231).  A scene whose next prop finds no free pose within
        ``max_attempts_per_prop`` raises for a strict (single-scene) env exactly like the reference; in a batch that scene alone
        starts its placement over (its earlier props happened to block the workspace), and is flagged ``failed`` if that keeps
        happening."""
        p = self._physics
        pcfg = self._cfg["prop_placer"]
        active = p.prop_active
        qadr = p.model.prop_qposadr
        for restart in range(self.PLACEMENT_RESTARTS):
            # contacts disabled for all props not placed yet (prop_initializer.py:183-185)
            active[mask] = 0
            self._park_missing_props(mask)
            stuck = torch.zeros_like(mask)
            for k in range(self.max_props):
                todo = mask & self.has_prop[:, k] & ~stuck
                if not bool(todo.any()):
                    continue
                active[todo, k] = 1
                if self._sampler == "device":
                    # one launch per prop: every scene's warp loops proposal -> kinematics -> collision pass -> rule on the device
                    w = self._cfg["task"]["initializers"]["workspace"]
                    lo = torch.tensor(w["min_pose"], device=self.device, dtype=self.dtype).expand(self.num_envs, 3)
                    hi = torch.tensor(w["max_pose"], device=self.device, dtype=self.dtype).expand(self.num_envs, 3)
                    self._place_seed += 1
                    out, quat = p.place_props(torch.full((self.num_envs,), k, device=self.device), lo, hi, self._table_gid, 0.6,
                                              int(pcfg["max_attempts_per_prop"]), self._place_seed, mask=todo)
                    ok = todo & (out[:, 3] > 0)
                    pose = torch.cat([out[:, :3], quat], 1)
                    p.qpos[:, qadr[k]:qadr[k] + 7] = torch.where(ok[:, None], pose, p.qpos[:, qadr[k]:qadr[k] + 7])
                    bad = todo & ~ok
                    if bool(bad.any()):
                        if self.strict:
                            raise RuntimeError("Failed to find a non-colliding pose for prop_%d after %d attempts" % (k, pcfg["max_attempts_per_prop"]))
                        stuck |= bad
                    continue
                for attempt in range(int(pcfg["max_attempts_per_prop"])):   # per-scene RandomState streams (reference draw order), one query launch per round
                    envs = torch.nonzero(todo).flatten()
                    pos, quat = self._sample_poses(envs.cpu().numpy())
                    pose = torch.tensor(np.concatenate([pos, quat], 1), device=self.device, dtype=self.dtype)
                    p.qpos[envs, qadr[k]:qadr[k] + 7] = pose
                    p.forward(full_contacts=True)
                    todo = todo & self._collides(k)
                    if not bool(todo.any()):
                        break
                else:
                    if self.strict:
                        raise RuntimeError("Failed to find a non-colliding pose for prop_%d after %d attempts" % (k, pcfg["max_attempts_per_prop"]))
                    stuck |= todo
            if not bool(stuck.any()):
                return
            # park every prop of the stuck scenes and start their placement over
            for k in range(self.max_props):
                p.qpos[stuck, qadr[k]:qadr[k] + 7] = torch.tensor([PARK_XY[0] + k, PARK_XY[1], 0.2, 1, 0, 0, 0], device=self.device, dtype=self.dtype)
            mask = stuck
        self.failed |= mask

    def _park_missing_props(self, mask: torch.Tensor):
        p = self._physics
        for k in range(self.max_props):
            m = mask & ~self.has_prop[:, k]
            if bool(m.any()):
                a = p.model.prop_qposadr[k]
                p.qpos[m, a:a + 7] = torch.tensor([PARK_XY[0] + k, PARK_XY[1], 0.2, 1, 0, 0, 0], device=self.device, dtype=self.dtype)

    def _place_and_settle(self, mask: torch.Tensor):
        """place_and_settle (prop_initializer.py:233-278): place, settle with the arm held static, re-place scenes that did
        not come to rest, at most max_settle_physics_attempts times."""
        p = self._physics
        pcfg = self._cfg["prop_placer"]
        h = p.model.opt.timestep
        max_steps = int(round(pcfg["max_settle_physics_time"] / h)); min_steps = int(round(pcfg["min_settle_physics_time"] / h))
        todo = mask.clone()
        for _ in range(int(pcfg["max_settle_physics_attempts"])):
            self._place_props(todo)
            t0 = p.time.clone()
            # a freshly placed prop has no velocity (set_pose writes qpos; mj_forward leaves qvel)
            p.qvel[todo, 15:] = 0
            done = p.settle(max_steps, min_steps, pcfg["max_qvel_tol"], pcfg["max_qacc_tol"], mask=todo)
            unsettled = todo & (done >= max_steps)
            if not bool(unsettled.any()):
                return True
            p.time[unsettled] = t0[unsettled]                      # prop_initializer.py:258
            todo = unsettled
        self.settle_failed = todo
        return False

    # ------------------------------------------------------------------ dm_env API
    def reset(self, mask: Optional[torch.Tensor] = None, resample_props: bool = False) -> TimeStep:
        """``mask`` resets a subset of scenes; ``resample_props`` re-draws their prop count / size / colour first (a new env in
        the reference)."""
        p = self._physics
        m = torch.ones(self.num_envs, dtype=torch.bool, device=self.device) if mask is None else mask.to(self.device, torch.bool)
        if resample_props:
            self.resample_scenes(m)
        p.reset(m)                                                                            # rearrangement.py:302
        home = torch.tensor(self._cfg["robots"]["arm"]["configuration"]["home"], device=self.device, dtype=self.dtype)
        p.qpos[m, 0:7] = home                                                                 # :306
        self.failed[m] = False
        self._park_missing_props(m)
        self._place_and_settle(m)                                                             # :309
        if self._robot is None:
            self._robot = BatchedRobotArm(p, self._osc, self._minmax, self._cfg["control_dt"])   # :319-324
        self._minmax.set_status(torch.zeros(self.num_envs, dtype=torch.uint8, device=self.device), m)
        p.forward(full_contacts=False)
        home_pose = self._robot.eef_pose.clone(); home_pose[:, 0] -= 0.1                       # :327-328
        if self.eef_home_pose is None or mask is None:
            self.eef_home_pose = home_pose
        else:
            self.eef_home_pose[m] = home_pose[m]
        p.status(clear=m)                                                                     # the reset scenes' flags only
        self.mode = "pick"                                                                    # :331
        return TimeStep(StepType.FIRST, 0.0, 0.0, self._compute_observation())

    def step(self, action_dict, active: Optional[torch.Tensor] = None) -> TimeStep:
        """``active`` (bool [n_env], batched variant only): scenes outside it are not stepped (their episode is over, as in the
        reference's data generator, which stops calling ``step`` on an env whose demonstration is complete or has failed)."""
        observation = self._compute_observation()                                             # :343 (observation BEFORE acting)
        if self.mode == "pick":
            self.pick(action_dict["pose"], active); self.mode = "place"
        else:
            self.place(action_dict["pose"], active); self.mode = "pick"
        return TimeStep(StepType.MID, 0.0, 0.0, observation)

    def _phase(self, duration: float, what: str):
        ok = self._robot.run_controller(duration)
        bad = ~ok if self._active is None else ~ok & self._active          # scenes outside the mask were not stepped
        self.failed |= bad
        if self.strict and bool(bad.any()):
            raise RuntimeError("Failed to " + what)                                           # rearrangement.py:372,377,...

    def _pose_tensor(self, pose) -> torch.Tensor:
        t = pose if isinstance(pose, torch.Tensor) else torch.as_tensor(np.asarray(pose))
        if t.dim() == 1:
            t = t.unsqueeze(0).expand(self.num_envs, 7)
        if isinstance(pose, (torch.Tensor, np.ndarray)):
            pose[..., 2] = 0.575                                                              # the reference mutates the caller's array (:362)
        t = t.to(self.device, self.dtype).clone(); t[:, 2] = 0.575
        return t

    def _pick_or_place(self, pose, grip_status: str, names, active: Optional[torch.Tensor] = None):
        osc, zeros = self._osc, torch.zeros(3, device=self.device, dtype=self.dtype)
        pose = self._pose_tensor(pose)
        pre = pose.clone(); pre[:, 2] = 0.9
        if active is not None:                      # idle scenes (episode over) are not stepped at all: b2p_set_env_mask
            self._active = active.to(self.device, torch.bool)
            self._physics.set_env_mask(self._active)
            idle = ~self._active
            home = torch.cat([self.eef_home_pose, torch.tensor(DOWN_QUAT, device=self.device, dtype=self.dtype).expand(self.num_envs, 4)], 1)
            pose[idle] = home[idle]; pre[idle] = home[idle]
        if grip_status == "max":
            osc.set_target(position=pre[:, :3], velocity=zeros, quat=pre[:, 3:], angular_velocity=zeros)    # :365-370
        else:
            osc.set_target(position=pre[:, :3], quat=pre[:, 3:])                                            # :411-414
        self._phase(2.0, names[0])
        osc.set_target(position=pose[:, :3]); self._phase(2.0, names[1])
        if active is None:
            self._minmax.status = grip_status
        else:
            self._minmax.set_status(torch.full((self.num_envs,), 1 if grip_status == "max" else 0, dtype=torch.uint8, device=self.device), ~idle)
        self._phase(1.0, names[2])
        osc.set_target(position=pre[:, :3]); self._phase(2.0, names[3])
        osc.set_target(position=self.eef_home_pose, quat=torch.tensor(DOWN_QUAT, device=self.device, dtype=self.dtype))
        self._phase(2.0, names[4])
        if active is not None:
            self._active = None
            self._physics.set_env_mask(None)

    def pick(self, pose, active: Optional[torch.Tensor] = None):
        """rearrangement.py:358-401."""
        self._pick_or_place(pose, "max", ("move arm to pre pick position", "move arm to pick position", "close gripper",
                                          "move arm to pre grasp position", "move arm to home position"), active)

    def place(self, pose, active: Optional[torch.Tensor] = None):
        """rearrangement.py:403-440."""
        self._pick_or_place(pose, "min", ("move arm to pre place position", "move arm to place position", "open gripper",
                                          "move arm to pre place position", "move arm to home position"), active)

    def random_pick_and_place(self):
        """rearrangement.py:667-698, per scene: pick pose = position of the scene's first prop + grasp quat
        mju_mat2Quat(R_xyz(0, 180, yaw) in degrees) with yaw = atan2(R10, R00) of the prop's geom_xmat, signed (``props_info`` stores
        ``R.from_matrix(geom_xmat).as_quat()``, i.e. xyzw, which ``R.from_quat`` reads back unchanged: rearrangement.py:246-247,676-679);
        place pose = U(workspace) + the same quat -> (pick [n, 7], place [n, 7])."""
        p = self._physics
        first = self.has_prop.to(torch.int64).argmax(1)
        e = torch.arange(self.num_envs, device=self.device)
        prop = torch.stack([p.qpos[:, a:a + 7] for a in p.model.prop_qposadr], 1)[e, first]
        # props_info is rebuilt from geom_xmat of the current state (rearrangement.py:240-258): one forward launch, then the kernel's own matrix
        p.forward(full_contacts=False)
        gid = torch.as_tensor(p.model.prop_geomid, device=self.device)[first]
        R = p.data.geom_xmat[e, gid]
        yaw = torch.atan2(R[:, 3], R[:, 0])
        zero = torch.zeros_like(yaw)
        quat = mat2quat(euler_xyz_matrix(zero, zero + math.pi, yaw)).to(self.dtype)
        ws = self._cfg["task"]["initializers"]["workspace"]
        lo, hi = (torch.tensor(ws[k], device=self.device, dtype=self.dtype) for k in ("min_pose", "max_pose"))
        place = lo + (hi - lo) * torch.rand(self.num_envs, 3, device=self.device, dtype=self.dtype, generator=self._dgen)
        return torch.cat([prop[:, :3].clone(), quat], 1), torch.cat([place, quat], 1)

    # ------------------------------------------------------------------ scripted expert + camera
    @property
    def expert(self) -> SortColoursExpert:
        if self._expert is None:
            self._expert = SortColoursExpert(self)
        return self._expert

    def sort_colours(self):
        """rearrangement.py:700-751 per scene -> (in_progress [n], pick_pose [n, 7], place_pose [n, 7])."""
        in_progress, pick, place, _ = self.expert.sort_colours()
        return in_progress, pick, place

    def prop_pick(self, prop_id):
        return self.expert.prop_pick(torch.as_tensor(prop_id, device=self.device).expand(self.num_envs).to(torch.int64))

    def prop_place(self, prop_id, min_pose=None, max_pose=None):
        w = self._cfg["task"]["initializers"]["workspace"]
        lo = torch.as_tensor(w["min_pose"] if min_pose is None else min_pose, device=self.device, dtype=self.dtype)
        hi = torch.as_tensor(w["max_pose"] if max_pose is None else max_pose, device=self.device, dtype=self.dtype)
        ex = lambda t: t.expand(self.num_envs, 3) if t.dim() == 1 else t
        return self.expert.prop_place(torch.as_tensor(prop_id, device=self.device).expand(self.num_envs).to(torch.int64), ex(lo), ex(hi))

    def render(self, lo: int = 0, hi: Optional[int] = None, seg: bool = False) -> Dict[str, torch.Tensor]:
        """``_compute_observation`` of the reference (rearrangement.py:460-478) for scenes [lo, hi): overhead rgb u8 [n, H, W, 3]
        and depth f32 [n, H, W] in metres (+ ``segmentation`` geom ids with ``seg``), ray-cast on the GPU."""
        if self._renderer is None:
            from .render import BatchedRenderer
            self._renderer = BatchedRenderer(self.device)
        o = self._renderer.render(self._physics, self.prop_rgba, lo=lo, hi=hi, seg=seg)
        obs = {"overhead_camera/rgb": o["rgb"], "overhead_camera/depth": o["depth"]}
        if seg:
            obs["segmentation"] = o["seg"]
        return obs

    def pixel_2_world(self, camera_name, coords):
        """rearrangement.py:507-535: pixel [n, 2] -> world [n, 3] through the rendered depth at that pixel."""
        c = torch.as_tensor(coords, device=self.device)
        d = self.render()["overhead_camera/depth"]
        r = torch.round(c).to(torch.int64)
        return self.camera.pixel_2_world(c, d[torch.arange(self.num_envs, device=self.device), r[:, 1], r[:, 0]])

    def world_2_pixel(self, camera_name, coords):
        return self.camera.world_2_pixel(coords)

    def get_camera_metadata(self):
        return self.camera.metadata()

    # ------------------------------------------------------------------ specs / observation
    def _compute_observation(self) -> Dict[str, torch.Tensor]:
        """State observation (views / small gathers of device tensors) and, with ``camera_obs``, the reference's observation keys
        ``overhead_camera/rgb`` / ``overhead_camera/depth`` (rearrangement.py:460-478) ray-cast on the GPU."""
        p = self._physics
        qa = p.model.prop_qposadr
        prop = torch.stack([p.qpos[:, a:a + 7] for a in qa], 1)
        # copies, not views: the TimeStep of step() is the observation BEFORE acting (rearrangement.py:343) and pick / place advance
        # the state buffers in place
        obs = {"qpos": p.qpos.clone(), "qvel": p.qvel.clone(), "time": p.time.clone(), "eef_pos": p.data.site_xpos[:, 1].clone(), "prop_pos": prop[..., :3],
               "prop_quat": prop[..., 3:], "has_prop": self.has_prop.clone()}
        if self.camera_obs:
            obs.update(self.render())
        return obs

    def observation_spec(self):
        p = self._physics
        spec = {"qpos": (self.num_envs, p.model.nq), "qvel": (self.num_envs, p.model.nv), "time": (self.num_envs,), "eef_pos": (self.num_envs, 3),
                "prop_pos": (self.num_envs, self.max_props, 3), "prop_quat": (self.num_envs, self.max_props, 4), "has_prop": (self.num_envs, self.max_props)}
        if self.camera_obs:       # rearrangement.py:442-451
            spec["overhead_camera/depth"] = (self.num_envs, self.overhead_camera_height, self.overhead_camera_width)
            spec["overhead_camera/rgb"] = (self.num_envs, self.overhead_camera_height, self.overhead_camera_width, 3)
        return spec

    def action_spec(self):
        """rearrangement.py:448-458: pose f64[7] (+ pixel_coords i64[2], gripper_rot f64[1] consumed by the data generator)."""
        return {"pose": (self.num_envs, 7), "pixel_coords": (self.num_envs, 2), "gripper_rot": (self.num_envs, 1)}

    @property
    def props_info(self):
        """rearrangement.py:277-295 (poses only; labels are render-side): per scene/prop position + orientation."""
        p = self._physics
        prop = torch.stack([p.qpos[:, a:a + 7] for a in p.model.prop_qposadr], 1)
        return {"position": prop[..., :3], "orientation": prop[..., 3:], "has_prop": self.has_prop, "half_size": self.prop_half_size,
                "rgba": self.prop_rgba, "colour": self.prop_colour}

    def prop_bboxes(self) -> torch.Tensor:
        """``get_bbox`` of ``props_info`` (rearrangement.py:251-283): PASCAL-VOC box (xmin, ymin, xmax, ymax) of every prop in the
        segmentation image of the overhead camera -> int64 [n_env, max_props, 4]; (-1, -1, -1, -1) for a prop out of view."""
        seg = self.render(seg=True)["segmentation"]
        n, H, W = seg.shape
        out = torch.full((n, self.max_props, 4), -1, dtype=torch.int64, device=self.device)
        xs, ys = torch.arange(W, device=self.device), torch.arange(H, device=self.device)
        for k in range(self.max_props):
            m = seg == int(self._prop_gid[k])
            cols, rows = m.any(1), m.any(2)
            seen = cols.any(1)
            big = torch.iinfo(torch.int64).max
            x0 = torch.where(cols, xs, big).min(1).values; x1 = torch.where(cols, xs, -1).max(1).values
            y0 = torch.where(rows, ys, big).min(1).values; y1 = torch.where(rows, ys, -1).max(1).values
            box = torch.stack([x0, y0, x1, y1], 1)
            out[:, k] = torch.where(seen[:, None], box, out[:, k])
        return out


def RearrangementEnv(viewer=None, cfg=None, **kw) -> BatchedRearrangementEnv:
    """Single-scene façade with the reference's constructor signature ``RearrangementEnv(viewer=None, cfg=DEFAULT_CONFIG)``."""
    return BatchedRearrangementEnv(viewer=viewer, cfg=cfg, num_envs=1, **kw)
