"""Batched counterparts of the reference's controller plug-ins (SURVEY.md 8b "plugin points").

Reference interfaces mirrored, with a leading scene dimension on every array:
  * ``MujocoController`` ABC  (mujoco_controllers/__init__.py:5-14): ``compute_control_output()``, ``is_converged()``;
  * ``OSC(physics, arm, controller_config)``  (mujoco_controllers/osc.py:33-97,153-157,180-247): ``set_target`` (sticky
    partial update, quaternion normalised), ``controller_gains`` (mutated in place by the tuner,
    automated_controller_tuning/rearrangement_controller_tuning.py:164-168), ``current_position_error()``,
    ``current_orientation_error()``;
  * ``MinMax(max_val, min_val)``  (mujoco_controllers/min_max.py:7-33): ``status`` in {"min", "max"}.

The arithmetic (Jacobian, task-space inertia, PD terms, null-space projection) is NOT done here: it runs fused with the
physics steps inside the sm_100a kernel (csrc/b2p_osc.h).  These classes hold the per-scene targets / gains as torch
tensors and hand them to the C-ABI; ``compute_control_output()`` is a zero-tick launch that evaluates the torque for
the current state (the reference's call order), used by callers that want the command without stepping.
"""
from __future__ import annotations

import abc
import copy
from dataclasses import dataclass
from typing import Optional

import torch

from .physics import ARM_HOME, BatchedPhysics

# config/robots/arm/controller_config/osc.yaml:1-22
DEFAULT_OSC_CONFIG = {
    "name": "osc", "physics_dt": 0.001, "control_dt": 0.005,
    "gains": {"position": {"kp": 350.0, "kd": 20.0}, "orientation": {"kp": 500.0, "kd": 100.0}, "nullspace": {"kp": 200.0, "kd": 30.0}},
    "nullspace": {"joint_config": list(ARM_HOME)},
    "convergence": {"position_threshold": 5e-3, "orientation_threshold": 68e-3},
}


class MujocoController(abc.ABC):
    """mujoco_controllers/__init__.py:5-14."""

    @abc.abstractmethod
    def compute_control_output(self):
        pass

    @abc.abstractmethod
    def is_converged(self):
        pass


@dataclass
class EEFTarget:
    """osc.py:26-31, one row per scene."""
    position: Optional[torch.Tensor]
    velocity: Optional[torch.Tensor]
    quat: Optional[torch.Tensor]
    angular_velocity: Optional[torch.Tensor]


def _gains6(g) -> list:
    return [float(g["position"]["kp"]), float(g["position"]["kd"]), float(g["orientation"]["kp"]), float(g["orientation"]["kd"]),
            float(g["nullspace"]["kp"]), float(g["nullspace"]["kd"])]


class BatchedOSC(MujocoController):
    """``mujoco_controllers.osc.OSC`` for ``physics.n_env`` scenes.  Same constructor shape as the reference's hydra partial
    (``controller(physics, arm)`` with ``controller_config`` bound, osc.yaml:24-27); ``arm`` is accepted and ignored (the arm's
    joints / attachment site are fixed by the compiled model)."""

    def __init__(self, physics: BatchedPhysics, arm=None, controller_config: Optional[dict] = None):
        self.physics = physics
        self.arm = arm
        cfg = copy.deepcopy(DEFAULT_OSC_CONFIG if controller_config is None else dict(controller_config))
        self.control_steps = int(cfg["control_dt"] // physics.model.opt.timestep)          # osc.py:47
        self.controller_gains = cfg["gains"]                                                # osc.py:48 (mutable, like the reference)
        self.position_threshold = float(cfg["convergence"]["position_threshold"])
        self.orientation_threshold = float(cfg["convergence"]["orientation_threshold"])
        self.nullspace_config = [float(x) for x in cfg["nullspace"]["joint_config"]]
        self.eef_target = EEFTarget(None, None, None, None)
        self._pushed = None
        kw = dict(device=physics.device, dtype=physics.dtype)
        self._target13 = torch.zeros(physics.n_env, 13, **kw)
        self._target13[:, 6] = 1.0
        self._grip = torch.zeros(physics.n_env, device=physics.device, dtype=torch.uint8)
        self._conv = torch.zeros(physics.n_env, device=physics.device, dtype=torch.bool)

    # ------------------------------------------------------------------ targets
    def _rows(self, x, width):
        t = torch.as_tensor(x, device=self.physics.device, dtype=self.physics.dtype)
        if t.dim() == 1:
            t = t.unsqueeze(0).expand(self.physics.n_env, width)
        assert t.shape == (self.physics.n_env, width), (tuple(t.shape), width)
        return t

    def set_target(self, position=None, velocity=None, quat=None, angular_velocity=None, mask: Optional[torch.Tensor] = None):
        """osc.py:86-97: fields left at None keep their previous value; the quaternion is normalised.  A 1-D argument is
        broadcast to every scene; ``mask`` (bool [n_env]) restricts the update to a subset of scenes."""
        def put(lo, width, val, normalise=False):
            if val is None:
                return
            t = self._rows(val, width)
            if normalise:
                t = t / t.norm(dim=1, keepdim=True)
            if mask is None:
                self._target13[:, lo:lo + width] = t
            else:
                self._target13[mask, lo:lo + width] = t[mask]
        put(0, 3, position); put(3, 3, velocity); put(6, 4, quat, True); put(10, 3, angular_velocity)
        t = self._target13
        self.eef_target = EEFTarget(position=t[:, 0:3], velocity=t[:, 3:6], quat=t[:, 6:10], angular_velocity=t[:, 10:13])

    @property
    def target13(self) -> torch.Tensor:
        return self._target13

    def set_gains_per_scene(self, gains: Optional[torch.Tensor]):
        """Batched-only: one gain vector per scene, [n_env, 6] = (position kp, kd, orientation kp, kd, nullspace kp, kd) -- a
        population of tuner candidates evaluated in one launch.  None returns to ``controller_gains`` for every scene."""
        self.physics.set_env_gains(gains)

    # ------------------------------------------------------------------ kernel hand-off
    def push_config(self):
        """Send gains / null-space posture / thresholds to the handle when they changed (the tuner mutates
        ``controller_gains`` between roll-outs)."""
        key = (tuple(_gains6(self.controller_gains)), tuple(self.nullspace_config), self.position_threshold, self.orientation_threshold)
        if key != self._pushed:
            self.physics.osc_config(key[0], key[1], key[2], key[3])
            self._pushed = key

    def compute_control_output(self, grip: Optional[torch.Tensor] = None) -> torch.Tensor:
        """osc.py:180-238 for the current state of every scene -> tau [n_env, 7] (zero-tick launch, no stepping)."""
        self.push_config()
        g = self._grip if grip is None else grip
        self._conv = self.physics.run_controller(self._target13, g, 0, self.control_steps).to(torch.bool).clone()
        return self.physics.derived("CTRL")[:, :7]

    def is_converged(self) -> torch.Tensor:
        """osc.py:240-247 evaluated by the kernel: bool [n_env] (latched over the last ``run_controller`` call)."""
        return self._conv

    # ------------------------------------------------------------------ diagnostics (tuner fitness; not on the hot path)
    def current_position_error(self) -> torch.Tensor:
        """osc.py:156-157: ||x* - x||_2 per scene (reads the derived site pose of the last launch)."""
        return (self._target13[:, 0:3] - self.physics.data.site_xpos[:, 0]).norm(dim=1)

    def current_orientation_error(self) -> torch.Tensor:
        """osc.py:153-154 / 132-150: sign(e_w) * e_xyz of  q* (x) conj(q)  per scene -> [n_env, 3]."""
        q = _mat2quat(self.physics.data.site_xmat[:, 0])
        qc = q * torch.tensor([1.0, -1.0, -1.0, -1.0], device=q.device, dtype=q.dtype)
        e = _mulquat(self._target13[:, 6:10], qc)
        return e[:, 1:4] * torch.sign(e[:, 0:1])


def _mulquat(a, b):
    aw, ax, ay, az = a.unbind(1); bw, bx, by, bz = b.unbind(1)
    return torch.stack([aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                        aw * by - ax * bz + ay * bw + az * bx, aw * bz + ax * by - ay * bx + az * bw], 1)


def _mat2quat(m):
    """mju_mat2Quat branch order, batched; m is [n, 9] row-major."""
    m00, m01, m02, m10, m11, m12, m20, m21, m22 = m.unbind(1)
    tr = m00 + m11 + m22
    one = torch.ones_like(tr)
    def nz(x):
        return torch.where(x.abs() < 1e-30, one * 1e-30, x)
    w0 = 0.5 * torch.sqrt((1 + tr).clamp_min(0)); q0 = torch.stack([w0, 0.25 * (m21 - m12) / nz(w0), 0.25 * (m02 - m20) / nz(w0), 0.25 * (m10 - m01) / nz(w0)], 1)
    x1 = 0.5 * torch.sqrt((1 + m00 - m11 - m22).clamp_min(0)); q1 = torch.stack([0.25 * (m21 - m12) / nz(x1), x1, 0.25 * (m01 + m10) / nz(x1), 0.25 * (m02 + m20) / nz(x1)], 1)
    y2 = 0.5 * torch.sqrt((1 - m00 + m11 - m22).clamp_min(0)); q2 = torch.stack([0.25 * (m02 - m20) / nz(y2), 0.25 * (m01 + m10) / nz(y2), y2, 0.25 * (m12 + m21) / nz(y2)], 1)
    z3 = 0.5 * torch.sqrt((1 - m00 - m11 + m22).clamp_min(0)); q3 = torch.stack([0.25 * (m10 - m01) / nz(z3), 0.25 * (m02 + m20) / nz(z3), 0.25 * (m12 + m21) / nz(z3), z3], 1)
    b0 = (tr > 0).unsqueeze(1); b1 = ((m00 > m11) & (m00 > m22)).unsqueeze(1); b2 = (m11 > m22).unsqueeze(1)
    q = torch.where(b0, q0, torch.where(b1, q1, torch.where(b2, q2, q3)))
    return q / q.norm(dim=1, keepdim=True)


class BatchedMinMax(MujocoController):
    """``mujoco_controllers.min_max.MinMax`` with one status bit per scene.  ``status = "max"`` / ``"min"`` sets every scene
    (the reference's setter, min_max.py:26-30); ``set_status(mask_or_bits)`` sets them individually."""

    def __init__(self, max_val: float = 255.0, min_val: float = 0.0, n_env: int = 1, device=None):
        self.max_val, self.min_val = float(max_val), float(min_val)
        self._bits = torch.zeros(n_env, dtype=torch.uint8, device=device)

    @property
    def status(self):
        on = int(self._bits.sum().item())
        return "max" if on == self._bits.numel() else ("min" if on == 0 else "mixed")

    @status.setter
    def status(self, new_status):
        assert new_status in ["min", "max"]
        self._bits.fill_(1 if new_status == "max" else 0)

    def set_status(self, bits: torch.Tensor, mask: Optional[torch.Tensor] = None):
        b = torch.as_tensor(bits, device=self._bits.device).to(torch.uint8)
        if mask is None:
            self._bits.copy_(b)
        else:
            self._bits[mask] = b[mask]

    @property
    def bits(self) -> torch.Tensor:
        return self._bits

    def compute_control_output(self) -> torch.Tensor:
        """min_max.py:14-19 -> [n_env] command values."""
        return torch.where(self._bits != 0, torch.full_like(self._bits, 0, dtype=torch.float64) + self.max_val,
                           torch.full_like(self._bits, 0, dtype=torch.float64) + self.min_val)

    def is_converged(self):
        pass
