"""BatchedLasaDrawEnv: the drawing task of the reference (``tasks/lasa_draw.py``) with a leading scene dimension.

The scene is the reference's: a 1 x 1 x 0.2 table, the Franka arm WITHOUT a gripper, a cylinder tool piece on the attachment body
(``lasa_draw.py:91-128``), compiled by ``compile/scene.py: build_draw_scene`` into ``assets/lasa_draw.b2pm`` (torque motors,
``generate_default_config``) and ``assets/lasa_draw_position.b2pm`` (position servos, ``generate_deployment_config``).  Without
the gripper's XML nothing sets ``<option cone>``, so these scenes run MuJoCo's default **pyramidal** friction cone; the kernels'
class-capable instantiations carry those rows (csrc/b2p_rows.h, ROW_PYR).

Mirrored methods (same names, argument meaning and return values, every argument with a leading ``[num_envs]`` dimension):

  reset(arm_configuration)                      lasa_draw.py:205-243
  interactive_tuning(mocap_pos, mocap_quat)     lasa_draw.py:300-324   one OSC tick towards the mocap body + (0, 0, 0.1)
  move_to_draw_target(position, velocity)       lasa_draw.py:326-349   one OSC tick with a velocity target -> (qpos, qvel, ctrl) of the arm
  move_to_joint_position_target(target)         lasa_draw.py:351-368   5 steps of the position servos (the reference's random push has scale 0)

``config/lasa.yaml`` sets ``physics_dt: 0.01`` and the ticks hold a command for 5 steps: with the OSC gains of
``controller_config/osc.yaml`` that sample-and-hold loop is not stable (orientation kd 100 x 0.05 s); ``physics_dt`` is therefore a
compile-time argument of ``compile_lasa_draw`` and the committed blobs keep the reference's value.

Observations (``_compute_observation``, lasa_draw.py:281-298: ``main_camera`` rgb + depth, 640 x 640): the reference's OpenGL
renderer is unavailable on this box; the frames come from the CUDA ray-caster (``render.BatchedRenderer``) on the scene's own render
model ``assets/lasa_draw.b2pr`` -- camera of config/arena/cameras/lasa.yaml, the arm's collision hulls, the tool piece as a 64-gon
prism, the translucent draw plane (lasa_draw.py:147-163) as a blended zone.
"""
from __future__ import annotations

from typing import Optional

import torch

from .physics import ARM_HOME, BatchedPhysics, load_default_blob
from .render import BatchedRenderer, load_default_render_blob

TOOL_OFFSET = (0.0, 0.0, 0.1)          # "account for height of toolpiece" (lasa_draw.py:312,334)
CONTROL_STEPS = 5                      # lasa_draw.py:320,343,361: for _ in range(5)


class BatchedLasaDrawEnv:
    def __init__(self, num_envs: int, dtype: torch.dtype = torch.float32, device: Optional[torch.device] = None, actuators: str = "motor"):
        if actuators not in ("motor", "position"):
            raise ValueError("actuators: 'motor' (generate_default_config) or 'position' (generate_deployment_config)")
        blob, names = load_default_blob("lasa_draw" if actuators == "motor" else "lasa_draw_position")
        self.actuators = actuators
        self.physics = BatchedPhysics(num_envs, dtype=dtype, device=device, blob=blob, names=names)
        self.num_envs, self.device, self.dtype = num_envs, self.physics.device, dtype
        self.physics.set_out_fields("SITE", "DYN", "GEOM")  # the eef site, the controller's torque command, the geom poses the renderer reads
        self._renderer = None
        self._grip = torch.zeros(num_envs, device=self.device, dtype=torch.uint8)      # no gripper: the OSC's gripper bit is ignored (nu = 7)
        self.mocap_quat = torch.tensor([0.0, 0.0, 1.0, 0.0], device=self.device, dtype=dtype).repeat(num_envs, 1)   # R.from_euler('xyz', [180, 180, 0]) as MuJoCo reads it (lasa_draw.py:137)
        self.reset()

    # ---------------------------------------------------------------------------------------------- lasa_draw.py:205-243
    def reset(self, arm_configuration=None):
        p = self.physics
        p.reset()
        q = torch.as_tensor(ARM_HOME if arm_configuration is None else arm_configuration, device=self.device, dtype=self.dtype)
        p.qpos[:, :7] = q
        p.forward(full_contacts=False)
        self.eef_home_pose = self.eef_pos.clone()
        return self.eef_home_pose

    @property
    def eef_pos(self) -> torch.Tensor:
        return self.physics.data.site_xpos[:, 0]          # the arm's attachment site (robot_arm.py:40: eef_site = arm.attachment_site)

    def _tick(self, position, quat, velocity):
        n = self.num_envs
        t = torch.zeros(n, 13, device=self.device, dtype=self.dtype)
        t[:, 0:3] = torch.as_tensor(position, device=self.device, dtype=self.dtype) + torch.tensor(TOOL_OFFSET, device=self.device, dtype=self.dtype)
        t[:, 3:6] = torch.as_tensor(velocity, device=self.device, dtype=self.dtype)
        t[:, 6:10] = torch.as_tensor(quat, device=self.device, dtype=self.dtype)
        self.physics.run_controller(t, self._grip, 1, CONTROL_STEPS)

    # ---------------------------------------------------------------------------------------------- lasa_draw.py:300-324
    def interactive_tuning(self, mocap_pos, mocap_quat=None):
        if self.actuators != "motor":
            raise RuntimeError("interactive_tuning drives the OSC: build the env with actuators='motor'")
        self._tick(mocap_pos, self.mocap_quat if mocap_quat is None else mocap_quat, torch.zeros(self.num_envs, 3, device=self.device, dtype=self.dtype))

    # ---------------------------------------------------------------------------------------------- lasa_draw.py:326-349
    def move_to_draw_target(self, target_position, target_velocity):
        if self.actuators != "motor":
            raise RuntimeError("move_to_draw_target drives the OSC: build the env with actuators='motor'")
        self._tick(target_position, self.mocap_quat, target_velocity)
        p = self.physics
        return p.qpos[:, :7].clone(), p.qvel[:, :7].clone(), p.derived("CTRL")[:, :7].clone()

    # ---------------------------------------------------------------------------------------------- lasa_draw.py:351-368
    def move_to_joint_position_target(self, target_position):
        if self.actuators != "position":
            raise RuntimeError("move_to_joint_position_target needs the position servos: build the env with actuators='position'")
        p = self.physics
        p.set_control(torch.as_tensor(target_position, device=self.device, dtype=self.dtype).expand(self.num_envs, 7))
        p.step(CONTROL_STEPS)          # xfrc_applied: the reference draws it with scale = 0 (lasa_draw.py:355-358), i.e. nothing is applied
        return p.qpos[:, :7].clone()

    # ---------------------------------------------------------------------------------------------- lasa_draw.py:245-298
    def observation_spec(self):
        r = self._get_renderer()
        return {"main_camera/depth": ((self.num_envs, r.height, r.width), torch.float32), "main_camera/rgb": ((self.num_envs, r.height, r.width, 3), torch.uint8)}

    def _get_renderer(self) -> BatchedRenderer:
        if self._renderer is None:
            self._renderer = BatchedRenderer(self.device, blob=load_default_render_blob("lasa_draw"))
        return self._renderer

    def _compute_observation(self):
        out = self._get_renderer().render(self.physics, seg=False)
        return {"main_camera/rgb": out["rgb"], "main_camera/depth": out["depth"]}

    def step(self, action_dict=None):
        """lasa_draw.py:245-257: the reference's step only observes (the motion goes through move_to_draw_target)."""
        return self._compute_observation()
