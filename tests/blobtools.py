"""Model-blob surgery for the analytic tests: unpack the committed B2PM blob, change options / arrays / counts, repack.
The result is a valid model of the same class (same tree, same geoms), so the oracle, the host build of the kernel templates
and the CUDA kernels all accept it.  Test infrastructure."""
import numpy as np

from mujoco_robot_environments_b200.compile import blob as B


def surgery(blob: bytes, opts=None, arrays=None, counts=None, edit=None) -> bytes:
    c, o, a = B.unpack(blob)
    if edit is not None:
        edit(c, o, a)
    o.update(opts or {})
    for k, v in (arrays or {}).items():
        a[k][...] = v
    c.update(counts or {})
    for name, dt, per, ckey in B.FIELDS:           # truncate arrays whose count was lowered (neq = 0, npair = 0)
        n = int(c.get(ckey, 0))
        if len(a[name]) > n:
            a[name] = a[name][:n]
    return B.pack(c, o, a)


def free_arm(blob: bytes, lock_all_but=None, timestep=None, integrator=None) -> bytes:
    """The arm + gripper as a frictionless, undamped, unconstrained tree under gravity: no joint damping / springs / limits, no
    equality constraints, no collision pairs, dead actuators.  `lock_all_but=d` adds an armature of 1e8 to every other tree dof
    (they stay where they are) and removes dof d's own armature -> a one-dof physical pendulum."""
    def edit(c, o, a):
        nt = c["narm"] + c["ngrip"]
        a["dof_damping"][:] = 0; a["jnt_stiffness"][:] = 0; a["jnt_limited"][:] = 0
        a["act_gain"][:] = 0; a["act_bias"][:] = 0; a["act_forcelimited"][:] = 0
        c["neq"] = 0; c["npair"] = 0
        if lock_all_but is not None:
            a["dof_armature"][:nt] = 1e8
            a["dof_armature"][lock_all_but] = 0
        if timestep is not None:
            o["timestep"] = timestep
        if integrator is not None:
            o["integrator"] = integrator
    return surgery(blob, edit=edit)


def tilted(blob: bytes, theta: float, mu: float, g: float = 9.8) -> bytes:
    """Gravity rotated by theta about the y axis (an incline of angle theta seen from the table) and sliding friction mu on the
    table and the props (both have priority 10, so the pair takes the element-wise maximum)."""
    def edit(c, o, a):
        o["gravity_x"], o["gravity_y"], o["gravity_z"] = g * np.sin(theta), 0.0, -g * np.cos(theta)
        for gi in [c["geom_table"]] + [int(x) for x in a["prop_geomid"]]:
            a["geom_friction"][gi, 0] = mu
    return surgery(blob, edit=edit)
