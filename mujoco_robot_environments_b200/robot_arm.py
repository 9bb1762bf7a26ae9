"""BatchedRobotArm: ``RobotArm`` (mujoco_robot_environments/models/robot_arm.py:17-94) for a batch of scenes.

``run_controller(duration)`` is THE hot path.  The reference runs, per scene and per control
tick, ``OSC.compute_control_output()`` + ``MinMax.compute_control_output()`` and then ``control_steps`` x
(``physics.set_control``, ``physics.step``), latching ``arm_converged``; with a gripper attached it never leaves the loop
early (robot_arm.py:65,86-88).  Here the whole loop -- every tick of every scene -- is ONE kernel launch through
``b2p_step_osc``.
"""
from __future__ import annotations

import math
from typing import Optional

import torch

from .controllers import BatchedMinMax, BatchedOSC
from .physics import BatchedPhysics

PINCH_SITE = "panda nohand/robotiq_2f85/pinch"          # robot_arm.py:57


class BatchedRobotArm:
    def __init__(self, physics: BatchedPhysics, arm_controller: Optional[BatchedOSC] = None,
                 end_effector_controller: Optional[BatchedMinMax] = None, control_dt: float = 0.005):
        self.physics = physics
        self.arm_controller = arm_controller if arm_controller is not None else BatchedOSC(physics)
        self.end_effector_controller = end_effector_controller if end_effector_controller is not None else \
            BatchedMinMax(255.0, 0.0, physics.n_env, physics.device)
        self.arm_joint_ids = list(range(7))                                                # robot_arm.py:40 (dofadr of joint1..7)
        self.end_effector_joint_ids = list(range(7, 15))
        self.control_steps = int(control_dt // physics.model.opt.timestep)                 # robot_arm.py:53
        self._tick_dt = self.control_steps * physics.model.opt.timestep

    @property
    def eef_pose(self) -> torch.Tensor:
        """robot_arm.py:55-58: world position of the pinch site, [n_env, 3] (derived output of the last launch)."""
        return self.physics.data.site_xpos[:, 1]

    def ticks_for(self, duration: float) -> int:
        """Number of passes of ``while time - start < duration`` (robot_arm.py:68) when every pass advances time by EXACTLY
        control_steps * timestep: 200 for 1.0 s, 400 for 2.0 s.

        Deviation, on purpose: the reference's ``data.time`` is a float64 accumulated 0.001 at a time, so ``time - start`` after 400
        passes is 2.0 -+ a few ulp, and for most start times the rounding falls on the low side -- the reference then runs a 401st
        (201st) pass, five more physics steps (0.25 % of the phase), depending on the scene's start time.  A batch runs one tick count
        for all scenes, so the exact-arithmetic count is used.  (``physics_env_steps`` of the data generator is
        sum(time) / h; in fp32 mode the record's time field is an fp32 accumulation, so that statistic is approximate.)"""
        return max(0, int(math.ceil(duration / self._tick_dt - 1e-9)))

    def run_controller(self, duration: float) -> torch.Tensor:
        """-> bool [n_env]: the reference's return value per scene (``arm_converged`` latched over the ticks)."""
        osc = self.arm_controller
        osc.push_config()
        conv = self.physics.run_controller(osc.target13, self.end_effector_controller.bits, self.ticks_for(duration), self.control_steps)
        osc._conv = conv.to(torch.bool).clone()
        return osc._conv
