"""Scene assembly + model compilation for the rearrangement model class.

``build_scene`` follows ``RearrangementEnv.__init__`` step by step (reference:
mujoco_robot_environments/tasks/rearrangement.py:54-210, call stack CS1 in
SURVEY.md): empty arena -> table -> arm (+gripper) -> props, with dm_control name
scoping.  ``compile_model`` plays the role of the MuJoCo compiler + ``mj_setConst``
for this model class and emits the flat arrays of ``blob.FIELDS``.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Sequence

import numpy as np

from . import blob as B
from .mjcf import (Body, Model, _floats, convex_hull, load_mesh, mat_to_quat, mesh_mass_properties,
                   quat_mul, quat_normalize, quat_to_mat)

DEFAULT_MENAGERIE = os.environ.get(
    "B2P_MENAGERIE",
    "/root/reference/mujoco_robot_environments/local_builds/madrona_mjx/external/mujoco_menagerie")

def _fmt(v):
    return " ".join(repr(float(x)) for x in v)


ARM_HOME = [0, -0.785, 0, -2.356, 0, 1.571, 0.785]  # config/robots/arm/franka_emika_panda.yaml:8


# ----------------------------------------------------------------------------- scene assembly
def _rectangle(name, half, mass, margin, gap) -> Model:
    """props.py:175-222 Rectangle._make: box geom condim 3, priority 10, friction (1,.005,.0001)."""
    m = Model(name)
    root = Body("prop_root", np.zeros(3), np.array([1.0, 0, 0, 0]), m.world)
    m.world.children.append(root)
    root.geoms.append(dict(name=name, type="box", size=_fmt(float(x) for x in half), condim="3",
                           mass=repr(float(mass)), friction="1 0.005 0.0001", margin=repr(float(margin)),
                           gap=repr(float(gap)), priority="10"))
    root.sites.append(dict(name="box_centre", pos="0 0 0"))
    return m


def build_scene(prop_half_sizes: Sequence[Sequence[float]], menagerie: str = DEFAULT_MENAGERIE,
                physics_dt=0.001, gravity=(0.0, 0.0, -9.8), prop_mass=0.1, prop_margin=0.15, prop_gap=0.15,
                arm_ctrlrange=((-87, 87),) * 4 + ((-12, 12),) * 3, gripper_forcerange=(-1.5, 1.5),
                arm_xml="franka_emika_panda/panda_nohand.xml", gripper_xml="robotiq_2f85/2f85.xml") -> Model:
    arena = Model("arena")
    arena.option.update(timestep=repr(physics_dt), gravity=_fmt(float(g) for g in gravity))
    # models/arenas/empty_assets/arena.xml:14-15
    arena.world.geoms.append(dict(name="ground", type="plane", size="2 2 2", friction="0.4"))
    # table (rearrangement.py:86-101)
    arena.world.sites.append(dict(name="table_center", pos="0.4 0.0 0.2"))
    arena.attach(_rectangle("table", (0.9, 1.0, 0.2), 10, 0.0, 0.0), "table_center")
    # arm: FER._build (models/arms/franka_emika.py:36-103): strip keyframes+actuators, add 7 motors
    arm = Model.from_path(os.path.join(menagerie, arm_xml))
    arm.actuators = []
    arm_joints = [j["name"] for j in arm.all_joints()]
    for i, jn in enumerate(arm_joints):
        lo, hi = arm_ctrlrange[i]
        arm.actuators.append(dict(type="motor", name="actuator%d" % (i + 1), joint=jn, ctrlrange=_fmt((lo, hi))))
    # gripper: Robotiq2F85._build (models/end_effectors/robotiq_2f85.py:35-51): forcerange override
    grip = Model.from_path(os.path.join(menagerie, gripper_xml))
    for a in grip.actuators:
        if a.get("name") == "fingers_actuator":
            a["forcerange"] = _fmt(float(x) for x in gripper_forcerange)
    arm.attach(grip, "attachment_site")  # standard_compose (models/robot_arm.py:97-112)
    arena.world.sites.append(dict(name="robot_base", pos="0.0 0.0 0.4"))
    arena.attach(arm, "robot_base")  # rearrangement.py:121-126
    # props (props.py:259-304, 583-639): free entities prop_i
    for i, hs in enumerate(prop_half_sizes):
        arena.attach(_rectangle("prop_%d" % i, hs, prop_mass, prop_margin, prop_gap), None, free=True)
    arena.nprop = len(prop_half_sizes)
    arena.autolimits = True
    return arena


# actuator_config/position.yaml: <general biastype="affine"> position servos, per joint (ctrlrange, forcerange, kp, kv)
POSITION_ACTUATORS = [("-2.8973 2.8973", "-87 87", 4500, 450), ("-1.7628 1.7628", "-87 87", 4500, 450), ("-2.8973 2.8973", "-87 87", 3500, 350),
                      ("-3.0718 -0.0698", "-87 87", 3500, 350), ("-2.8973 2.8973", "-12 12", 2000, 200), ("-0.0175 3.7525", "-12 12", 2000, 200),
                      ("-2.8973 2.8973", "-12 12", 2000, 200)]


def build_draw_scene(menagerie: str = DEFAULT_MENAGERIE, physics_dt=0.01, gravity=(0.0, 0.0, -9.8), actuators="motor",
                     arm_ctrlrange=((-87, 87),) * 4 + ((-12, 12),) * 3, arm_xml="franka_emika_panda/panda_nohand.xml") -> Model:
    """LasaDrawEnv.__init__ step by step (reference tasks/lasa_draw.py:60-176; config/lasa.yaml: physics_dt 0.01): empty arena ->
    1 x 1 x 0.2 table at (0.4, 0, 0.2) -> the arm WITHOUT a gripper, a cylinder tool piece (radius 0.015, half height 0.05) on the
    attachment body -> base at (0, 0, 0.4).  Without the gripper's XML nothing sets ``cone``, so the scene runs MuJoCo's default
    pyramidal cone.  The mocap target and the draw plane have contype = conaffinity = 0 and no mass: they do not enter the physics
    and are left out.  ``actuators``: "motor" (generate_default_config) or "position" (generate_deployment_config:
    robots/arm/actuator_config=position)."""
    arena = Model("arena")
    arena.option.update(timestep=repr(physics_dt), gravity=_fmt(float(g) for g in gravity))
    arena.world.geoms.append(dict(name="ground", type="plane", size="2 2 2", friction="0.4"))
    arena.world.sites.append(dict(name="table_center", pos="0.4 0.0 0.2"))
    arena.attach(_rectangle("table", (1.0, 1.0, 0.2), 10, 0.0, 0.0), "table_center")          # lasa_draw.py:91-108
    arm = Model.from_path(os.path.join(menagerie, arm_xml))
    arm.actuators = []
    for i, j in enumerate(arm.all_joints()):                                                   # franka_emika.py:66-103
        if actuators == "motor":
            lo, hi = arm_ctrlrange[i]
            arm.actuators.append(dict(type="motor", name="actuator%d" % (i + 1), joint=j["name"], ctrlrange=_fmt((lo, hi))))
        else:
            cr, fr, kp, kv = POSITION_ACTUATORS[i]
            arm.actuators.append(dict(type="general", name="actuator%d" % (i + 1), joint=j["name"], dyntype="none", biastype="affine",
                                      ctrlrange=cr, forcerange=fr, gainprm=repr(float(kp)), biasprm="0 %r %r" % (-float(kp), -float(kv))))
    body, _ = arm.find_site("attachment_site")
    body.geoms.append(dict(name="tool", type="cylinder", size="0.015 0.05", pos="0 0 0.05"))   # lasa_draw.py:113-120 (MuJoCo's geom defaults)
    arena.world.sites.append(dict(name="robot_base", pos="0.0 0.0 0.4"))
    arena.attach(arm, "robot_base")
    arena.nprop = 0
    arena.autolimits = True
    return arena


# ----------------------------------------------------------------------------- compiler
_GEOM_TYPES = {"plane": B.GEOM_PLANE, "sphere": B.GEOM_SPHERE, "cylinder": B.GEOM_CYLINDER,
               "box": B.GEOM_BOX, "mesh": B.GEOM_MESH}


def _hull_centroid(v):
    from scipy.spatial import ConvexHull
    h = ConvexHull(v)
    c0 = v[h.vertices].mean(0)
    a, b, c = (v[h.simplices[:, k]] - c0 for k in range(3))
    vol = np.abs(np.einsum("ij,ij->i", a, np.cross(b, c))) / 6
    cen = (a + b + c) / 4 + c0
    return (cen * vol[:, None]).sum(0) / vol.sum()


class _MeshCache:
    def __init__(self, model: Model):
        self.model, self.cache = model, {}

    def get(self, name):
        if name not in self.cache:
            info = self.model.meshes[name]
            v, f = load_mesh(info["file"])
            v = v * info["scale"]
            vol, com, inertia = mesh_mass_properties(v, f)
            hv, nbr = convex_hull(v)
            cen = _hull_centroid(v)
            self.cache[name] = dict(vol=vol, com=com, inertia=inertia, hull=hv, nbr=nbr, centroid=cen)
        return self.cache[name]


def _diag_inertia(I):
    """Symmetric 3x3 -> (principal moments, quaternion of principal frame)."""
    w, V = np.linalg.eigh(I)
    order = np.argsort(-w)  # descending, as mju_eig3
    w, V = w[order], V[:, order]
    if np.linalg.det(V) < 0:
        V[:, 2] = -V[:, 2]
    return w, mat_to_quat(V)


def _body_inertial(b: Body, meshes: _MeshCache):
    """-> (mass, ipos, iquat, inertia[3]) in the body frame (MuJoCo: explicit <inertial> wins,
    else accumulated from ALL geoms incl. visual ones, default density 1000)."""
    if b.inertial is not None:
        a = b.inertial
        mass = float(a["mass"])
        ipos = _floats(a.get("pos"), 3, [0, 0, 0])
        if "fullinertia" in a:
            f = _floats(a["fullinertia"])
            I = np.array([[f[0], f[3], f[4]], [f[3], f[1], f[5]], [f[4], f[5], f[2]]])
            w, q = _diag_inertia(I)
            return mass, ipos, q, w
        return mass, ipos, quat_normalize(_floats(a.get("quat"), 4, [1, 0, 0, 0])), _floats(a["diaginertia"])
    mass, mcom, parts = 0.0, np.zeros(3), []
    for g in b.geoms:
        gpos = _floats(g.get("pos"), 3, [0, 0, 0])
        R = quat_to_mat(quat_normalize(_floats(g.get("quat"), 4, [1, 0, 0, 0])))
        gtype = g.get("type", "mesh" if g.get("mesh") else "sphere")
        if gtype == "box":
            s = _floats(g["size"])
            vol = 8 * s[0] * s[1] * s[2]
            gm = float(g["mass"]) if "mass" in g else float(g.get("density", 1000)) * vol
            Ig = gm / 3 * np.diag([s[1] ** 2 + s[2] ** 2, s[0] ** 2 + s[2] ** 2, s[0] ** 2 + s[1] ** 2])
            c = gpos
        elif gtype == "mesh":
            mi = meshes.get(g["mesh"])
            gm = float(g["mass"]) if "mass" in g else float(g.get("density", 1000)) * mi["vol"]
            Ig = R @ (mi["inertia"] * gm / mi["vol"]) @ R.T
            c = gpos + R @ mi["com"]
        elif gtype == "cylinder":   # size = (radius, half height), axis z
            s = _floats(g["size"])
            vol = np.pi * s[0] ** 2 * 2 * s[1]
            gm = float(g["mass"]) if "mass" in g else float(g.get("density", 1000)) * vol
            Ig = gm * np.diag([s[0] ** 2 / 4 + s[1] ** 2 / 3, s[0] ** 2 / 4 + s[1] ** 2 / 3, s[0] ** 2 / 2])
            c = gpos
        elif gtype == "plane":
            continue
        else:
            raise NotImplementedError("inertia of geom type " + gtype)
        if gtype in ("box", "cylinder"):
            Ig = R @ Ig @ R.T
        if gm <= 0:
            continue
        parts.append((gm, c, Ig))
        mass += gm
        mcom += gm * c
    if mass <= 0:
        return 0.0, np.zeros(3), np.array([1.0, 0, 0, 0]), np.zeros(3)
    mcom /= mass
    I = np.zeros((3, 3))
    for gm, c, Ig in parts:
        d = c - mcom
        I += Ig + gm * (d @ d * np.eye(3) - np.outer(d, d))
    w, q = _diag_inertia(I)
    return mass, mcom, q, w


def fuse_static(model: Model, meshes: _MeshCache):
    """Merge every joint-less body into its parent (dynamically equivalent; MuJoCo's
    ``fusestatic``).  Inertials are made explicit first so that nothing is lost."""
    for b in model.bodies():
        if b is model.world:
            continue
        mass, ipos, iquat, inertia = _body_inertial(b, meshes)
        b.inertial = dict(mass=repr(float(mass)), pos=_fmt(ipos), quat=_fmt(iquat),
                          diaginertia=_fmt(inertia))

    def rec(b: Body):
        for c in list(b.children):
            rec(c)
        for c in list(b.children):
            if c.joints:
                continue
            R = quat_to_mat(c.quat)
            # geoms / sites / grandchildren re-expressed in b's frame
            for coll in (c.geoms, c.sites):
                for e in coll:
                    p = _floats(e.get("pos"), 3, [0, 0, 0])
                    q = quat_normalize(_floats(e.get("quat"), 4, [1, 0, 0, 0]))
                    e["pos"] = _fmt(c.pos + R @ p)
                    e["quat"] = _fmt(quat_mul(c.quat, q))
            b.geoms += c.geoms
            b.sites += c.sites
            for gc in c.children:
                gc.pos = c.pos + R @ gc.pos
                gc.quat = quat_mul(c.quat, gc.quat)
                gc.parent = b
            idx = b.children.index(c)
            b.children[idx:idx + 1] = c.children
            # inertia
            if b is not model.world:
                mc = float(c.inertial["mass"])
                if mc > 0:
                    mb = float(b.inertial["mass"])
                    pb, qb, ib = _floats(b.inertial["pos"]), _floats(b.inertial["quat"]), _floats(b.inertial["diaginertia"])
                    pc = c.pos + R @ _floats(c.inertial["pos"])
                    qc = quat_mul(c.quat, _floats(c.inertial["quat"]))
                    ic = _floats(c.inertial["diaginertia"])
                    m = mb + mc
                    com = (mb * pb + mc * pc) / m
                    I = np.zeros((3, 3))
                    for mm, p, q, ii in ((mb, pb, qb, ib), (mc, pc, qc, ic)):
                        Rm = quat_to_mat(q)
                        d = p - com
                        I += Rm @ np.diag(ii) @ Rm.T + mm * (d @ d * np.eye(3) - np.outer(d, d))
                    w, q = _diag_inertia(I)
                    b.inertial = dict(mass=repr(float(m)), pos=_fmt(com), quat=_fmt(q),
                                      diaginertia=_fmt(w))
    rec(model.world)


def _name_everything(model: Model):
    k = 0
    for b in model.bodies():
        for g in b.geoms:
            if not g.get("name"):
                g["name"] = "%s#geom%d" % (b.name, k)
            k += 1


def _flatten(model: Model, meshes: _MeshCache, geom_order: Optional[List[str]] = None):
    """Body tree -> MuJoCo-ordered flat arrays (depth-first bodies; joints/dofs/geoms in body order)."""
    bodies = model.bodies()
    bid = {id(b): i for i, b in enumerate(bodies)}
    A: Dict[str, list] = {n: [] for n, *_ in B.FIELDS}
    names = dict(body=[b.name for b in bodies], jnt=[], geom=[], site=[])
    nq = nv = 0
    geoms = []
    for i, b in enumerate(bodies):
        par = bid[id(b.parent)] if b.parent is not None else 0
        A["body_parentid"].append(par)
        A["body_pos"].append(b.pos)
        A["body_quat"].append(b.quat)
        mass, ipos, iquat, inertia = (0.0, np.zeros(3), np.array([1.0, 0, 0, 0]), np.zeros(3)) if i == 0 else _body_inertial(b, meshes)
        A["body_mass"].append(mass); A["body_ipos"].append(ipos); A["body_iquat"].append(iquat); A["body_inertia"].append(inertia)
        A["body_jntnum"].append(len(b.joints)); A["body_jntadr"].append(len(names["jnt"]) if b.joints else -1)
        dof0 = nv
        for j in b.joints:
            jt = j.get("type", "hinge")
            names["jnt"].append(j.get("name", ""))
            jid = len(names["jnt"]) - 1
            A["jnt_bodyid"].append(i); A["jnt_qposadr"].append(nq); A["jnt_dofadr"].append(nv)
            A["jnt_pos"].append(_floats(j.get("pos"), 3, [0, 0, 0]))
            if jt == "free":
                A["jnt_type"].append(B.JNT_FREE)
                A["jnt_axis"].append([0, 0, 1]); A["jnt_stiffness"].append(0.0)
                A["jnt_limited"].append(0); A["jnt_range"].append([0, 0]); A["jnt_margin"].append(0.0)
                A["jnt_solref"].append([0.02, 1]); A["jnt_solimp"].append([0.9, 0.95, 0.001, 0.5, 2])
                A["qpos0"] += list(b.pos) + list(b.quat); A["qpos_spring"] += list(b.pos) + list(b.quat)
                nd, nqj = 6, 7
            elif jt == "hinge":
                A["jnt_type"].append(B.JNT_HINGE)
                ax = _floats(j.get("axis"), 3, [0, 0, 1])
                A["jnt_axis"].append(ax / np.linalg.norm(ax))
                A["jnt_stiffness"].append(float(j.get("stiffness", 0)))
                rng = _floats(j.get("range"), 2, [0, 0])
                lim = j.get("limited", "auto")
                limited = (lim == "true") or (lim == "auto" and model.autolimits and "range" in j)
                A["jnt_limited"].append(int(limited)); A["jnt_range"].append(rng)
                A["jnt_margin"].append(float(j.get("margin", 0)))
                A["jnt_solref"].append(_floats(j.get("solreflimit"), 2, [0.02, 1]))
                A["jnt_solimp"].append(_floats(j.get("solimplimit"), 5, [0.9, 0.95, 0.001, 0.5, 2]))
                A["qpos0"].append(float(j.get("ref", 0))); A["qpos_spring"].append(float(j.get("springref", 0)))
                nd, nqj = 1, 1
            else:
                raise NotImplementedError("joint type " + jt)
            for k in range(nd):
                A["dof_bodyid"].append(i); A["dof_jntid"].append(jid)
                A["dof_armature"].append(float(j.get("armature", 0))); A["dof_damping"].append(float(j.get("damping", 0)))
            nq += nqj; nv += nd
        A["body_dofnum"].append(nv - dof0); A["body_dofadr"].append(dof0 if nv > dof0 else -1)
        for g in b.geoms:
            if int(g.get("contype", 1)) == 0 and int(g.get("conaffinity", 1)) == 0:
                continue
            geoms.append((i, g))
        for s in b.sites:
            names["site"].append(s.get("name", ""))
            A["site_bodyid"].append(i)
            A["site_pos"].append(_floats(s.get("pos"), 3, [0, 0, 0]))
            A["site_quat"].append(quat_normalize(_floats(s.get("quat"), 4, [1, 0, 0, 0])))
    nbody = len(bodies)
    # dof parents, weld / root ids
    last_dof = [-1] * nbody
    for i in range(nbody):
        par = A["body_parentid"][i]
        inherited = last_dof[par] if i > 0 else -1
        d0, dn = A["body_dofadr"][i], A["body_dofnum"][i]
        for k in range(dn):
            A["dof_parentid"].append(inherited if k == 0 else d0 + k - 1)
        last_dof[i] = d0 + dn - 1 if dn else inherited
    for i in range(nbody):
        par = A["body_parentid"][i]
        A["body_weldid"].append(i if (i == 0 or A["body_jntnum"][i] > 0) else A["body_weldid"][par])
        A["body_rootid"].append(0 if i == 0 else (i if par == 0 else A["body_rootid"][par]))
    # geoms (optionally in a prescribed order so fused / unfused models share geom ids)
    if geom_order is not None:
        geoms.sort(key=lambda bg: geom_order.index(bg[1]["name"]))
    mesh_ids: Dict[str, int] = {}
    nbr_total = 0
    for bi, g in geoms:
        names["geom"].append(g["name"])
        gtype = g.get("type", "mesh" if g.get("mesh") else "sphere")
        gpos = _floats(g.get("pos"), 3, [0, 0, 0])
        gquat = quat_normalize(_floats(g.get("quat"), 4, [1, 0, 0, 0]))
        size = _floats(g.get("size"), 3, [0, 0, 0])
        meshid = -1
        if gtype == "mesh":
            mname = g["mesh"]
            mi = meshes.get(mname)
            if mname not in mesh_ids:
                mesh_ids[mname] = len(mesh_ids)
                hv = mi["hull"] - mi["centroid"]
                A["mesh_vertadr"].append(len(A["mesh_vert"])); A["mesh_vertnum"].append(len(hv))
                base = len(A["mesh_vert"])
                for k, v in enumerate(hv):
                    A["mesh_vert"].append(v)
                    A["mesh_nbradr"].append(nbr_total); A["mesh_nbrnum"].append(len(mi["nbr"][k]))
                    A["mesh_nbr"] += [int(x) for x in mi["nbr"][k]]
                    nbr_total += len(mi["nbr"][k])
            meshid = mesh_ids[mname]
            gpos = gpos + quat_to_mat(gquat) @ mi["centroid"]
            hv = mi["hull"] - mi["centroid"]
            lo, hi = hv.min(0), hv.max(0)
            size = (hi - lo) / 2
            aabb = np.concatenate([(lo + hi) / 2, (hi - lo) / 2])
            rbound = np.linalg.norm(hv, axis=1).max()
        elif gtype == "box":
            aabb = np.concatenate([np.zeros(3), size]); rbound = np.linalg.norm(size)
        elif gtype == "plane":
            aabb = np.zeros(6); rbound = 0.0
        elif gtype == "cylinder":   # size = (radius, half height, 0) as in MuJoCo
            half = np.array([size[0], size[0], size[1]])
            aabb = np.concatenate([np.zeros(3), half]); rbound = np.hypot(size[0], size[1])
        else:
            raise NotImplementedError("collision geom type " + gtype)
        A["geom_type"].append(_GEOM_TYPES[gtype]); A["geom_bodyid"].append(bi)
        A["geom_condim"].append(int(g.get("condim", 3))); A["geom_priority"].append(int(g.get("priority", 0)))
        A["geom_meshid"].append(meshid); A["geom_size"].append(size); A["geom_pos"].append(gpos); A["geom_quat"].append(gquat)
        A["geom_friction"].append(_floats(g.get("friction"), 3, [1, 0.005, 0.0001]))
        A["geom_solref"].append(_floats(g.get("solref"), 2, [0.02, 1]))
        A["geom_solimp"].append(_floats(g.get("solimp"), 5, [0.9, 0.95, 0.001, 0.5, 2]))
        A["geom_solmix"].append(float(g.get("solmix", 1))); A["geom_margin"].append(float(g.get("margin", 0)))
        A["geom_gap"].append(float(g.get("gap", 0))); A["geom_rbound"].append(rbound); A["geom_aabb"].append(aabb)
    counts = dict(nq=nq, nv=nv, nbody=nbody, njnt=len(names["jnt"]), ngeom=len(geoms), nsite=len(names["site"]),
                  nmesh=len(mesh_ids), nmeshvert=len(A["mesh_vert"]), nmeshnbr=nbr_total)
    return A, counts, names


# ---- numpy FK + mass matrix at a configuration (compile-time only: mj_setConst, connect anchors)
def _fk(A, counts, qpos):
    nb = counts["nbody"]
    xpos, xquat = np.zeros((nb, 3)), np.zeros((nb, 4)); xquat[0, 0] = 1
    xanchor, xaxis = np.zeros((counts["njnt"], 3)), np.zeros((counts["njnt"], 3))
    for i in range(1, nb):
        p = A["body_parentid"][i]
        R = quat_to_mat(xquat[p])
        pos = xpos[p] + R @ np.asarray(A["body_pos"][i]); quat = quat_mul(xquat[p], np.asarray(A["body_quat"][i]))
        for k in range(A["body_jntnum"][i]):
            j = A["body_jntadr"][i] + k
            qa = A["jnt_qposadr"][j]
            if A["jnt_type"][j] == B.JNT_FREE:
                pos, quat = qpos[qa:qa + 3].copy(), quat_normalize(qpos[qa + 3:qa + 7])
                xanchor[j], xaxis[j] = pos, [0, 0, 1]
            else:
                Rq = quat_to_mat(quat)
                xanchor[j] = pos + Rq @ np.asarray(A["jnt_pos"][j]); xaxis[j] = Rq @ np.asarray(A["jnt_axis"][j])
                ang = qpos[qa] - A["qpos0"][qa]
                ax = np.asarray(A["jnt_axis"][j])
                quat = quat_mul(quat, np.concatenate([[np.cos(ang / 2)], np.sin(ang / 2) * ax]))
                pos = xanchor[j] - quat_to_mat(quat) @ np.asarray(A["jnt_pos"][j])
        xpos[i], xquat[i] = pos, quat
    return xpos, xquat, xanchor, xaxis


def _jac(A, counts, xanchor, xaxis, xquat, body, point):
    nv = counts["nv"]
    jp, jr = np.zeros((3, nv)), np.zeros((3, nv))
    b = body
    while b > 0:
        for k in range(A["body_jntnum"][b]):
            j = A["body_jntadr"][b] + k
            d = A["jnt_dofadr"][j]
            if A["jnt_type"][j] == B.JNT_FREE:
                jp[:, d:d + 3] = np.eye(3)
                R = quat_to_mat(xquat[b])
                for a in range(3):
                    jr[:, d + 3 + a] = R[:, a]
                    jp[:, d + 3 + a] = np.cross(R[:, a], point - xanchor[j])
            else:
                jr[:, d] = xaxis[j]; jp[:, d] = np.cross(xaxis[j], point - xanchor[j])
        b = A["body_parentid"][b]
    return jp, jr


def mass_matrix(A, counts, qpos):
    xpos, xquat, xanchor, xaxis = _fk(A, counts, qpos)
    nv = counts["nv"]
    M = np.diag(np.asarray(A["dof_armature"], float))
    for b in range(1, counts["nbody"]):
        m = A["body_mass"][b]
        if m <= 0:
            continue
        R = quat_to_mat(xquat[b])
        xipos = xpos[b] + R @ np.asarray(A["body_ipos"][b])
        Ri = R @ quat_to_mat(np.asarray(A["body_iquat"][b]))
        Iw = Ri @ np.diag(A["body_inertia"][b]) @ Ri.T
        jp, jr = _jac(A, counts, xanchor, xaxis, xquat, b, xipos)
        M += m * jp.T @ jp + jr.T @ Iw @ jr
    return M, (xpos, xquat, xanchor, xaxis)


def _set_const(A, counts):
    """mj_setConst subset: dof_invweight0, body_invweight0, meaninertia at qpos0."""
    q0 = np.asarray(A["qpos0"], float)
    M, (xpos, xquat, xanchor, xaxis) = mass_matrix(A, counts, q0)
    Minv = np.linalg.inv(M)
    nv = counts["nv"]
    dinv = np.diag(Minv).copy()
    for j in range(counts["njnt"]):
        if A["jnt_type"][j] == B.JNT_FREE:
            d = A["jnt_dofadr"][j]
            dinv[d:d + 3] = dinv[d:d + 3].mean(); dinv[d + 3:d + 6] = dinv[d + 3:d + 6].mean()
    A["dof_invweight0"] = list(dinv)
    biw = np.zeros((counts["nbody"], 2))
    for b in range(1, counts["nbody"]):
        if A["body_weldid"][b] == 0:
            continue
        R = quat_to_mat(xquat[b])
        xipos = xpos[b] + R @ np.asarray(A["body_ipos"][b])
        jp, jr = _jac(A, counts, xanchor, xaxis, xquat, b, xipos)
        biw[b, 0] = np.trace(jp @ Minv @ jp.T) / 3; biw[b, 1] = np.trace(jr @ Minv @ jr.T) / 3
    A["body_invweight0"] = [list(x) for x in biw]
    return float(np.trace(M) / nv), (xpos, xquat)


def compile_model(model: Model, fuse: bool = True):
    """-> (counts, opts, arrays, names).  ``fuse=False`` keeps MuJoCo's body list (welded
    bodies separate); ``fuse=True`` merges them (what the CUDA kernels run on)."""
    meshes = _MeshCache(model)
    _name_everything(model)
    U, cu, nu_ = _flatten(model, meshes)
    meaninertia, (xpos0, xquat0) = _set_const(U, cu)
    ubody = {n: i for i, n in enumerate(nu_["body"])}
    ujnt = {n: i for i, n in enumerate(nu_["jnt"])}
    geom_names = list(nu_["geom"])
    # static pair filter on the UNFUSED model
    excl = set()
    for a, b in model.excludes:
        excl.add((ubody[a], ubody[b])); excl.add((ubody[b], ubody[a]))
    pairs = []
    for g1 in range(cu["ngeom"]):
        for g2 in range(g1 + 1, cu["ngeom"]):
            b1, b2 = U["geom_bodyid"][g1], U["geom_bodyid"][g2]
            w1, w2 = U["body_weldid"][b1], U["body_weldid"][b2]
            if w1 == w2:
                continue  # same weld body (covers static-static: both weld to world)
            if w1 != 0 and w2 != 0:
                if U["body_weldid"][U["body_parentid"][w1]] == w2 or U["body_weldid"][U["body_parentid"][w2]] == w1:
                    continue
            if (b1, b2) in excl:
                continue
            # MuJoCo orders a pair by geom type (mj_collideGeoms swaps so type1 <= type2)
            pairs.append((g1, g2) if U["geom_type"][g1] <= U["geom_type"][g2] else (g2, g1))
    # connect anchors in body2 frame at qpos0
    eqs = []
    for e in model.equalities:
        data = np.zeros(11)
        if e["type"] == "connect":
            b1, b2 = ubody[e["body1"]], ubody[e["body2"]]
            anchor = _floats(e.get("anchor"), 3, [0, 0, 0])
            g = xpos0[b1] + quat_to_mat(xquat0[b1]) @ anchor
            data[0:3] = anchor; data[3:6] = quat_to_mat(xquat0[b2]).T @ (g - xpos0[b2])
            eqs.append((B.EQ_CONNECT, e["body1"], e["body2"], data, e))
        elif e["type"] == "joint":
            data[0:5] = _floats(e.get("polycoef"), 5, [0, 1, 0, 0, 0])
            eqs.append((B.EQ_JOINT, e["joint1"], e["joint2"], data, e))
        else:
            raise NotImplementedError("equality " + e["type"])
    if fuse:
        fuse_static(model, meshes)
        A, counts, names = _flatten(model, meshes, geom_order=geom_names)
        A["dof_invweight0"] = U["dof_invweight0"]
        A["body_invweight0"] = [U["body_invweight0"][ubody[n]] for n in names["body"]]
    else:
        A, counts, names = U, cu, nu_
    body_id = {n: i for i, n in enumerate(names["body"])}
    jnt_id = {n: i for i, n in enumerate(names["jnt"])}
    assert names["geom"] == geom_names and names["jnt"] == nu_["jnt"]
    A["geom_invweight0"] = [U["body_invweight0"][U["geom_bodyid"][g]] for g in range(cu["ngeom"])]
    A["pair_geom1"] = [p[0] for p in pairs]; A["pair_geom2"] = [p[1] for p in pairs]
    for t, o1, o2, data, e in eqs:
        A["eq_type"].append(t)
        ids = body_id if t == B.EQ_CONNECT else jnt_id
        A["eq_obj1id"].append(ids[o1]); A["eq_obj2id"].append(ids[o2]); A["eq_data"].append(data)
        A["eq_solref"].append(_floats(e.get("solref"), 2, [0.02, 1]))
        A["eq_solimp"].append(_floats(e.get("solimp"), 5, [0.9, 0.95, 0.001, 0.5, 2]))
    # actuators
    tendons = {t["name"]: t for t in model.tendons}
    for a in model.actuators:
        if a.get("joint"):
            dofs, coefs = [A["jnt_dofadr"][jnt_id[a["joint"]]], -1], [1.0, 0.0]
            n = 1
        else:
            t = tendons[a["tendon"]]
            assert len(t["joints"]) == 2
            dofs = [A["jnt_dofadr"][jnt_id[j]] for j, _ in t["joints"]]; coefs = [c for _, c in t["joints"]]
            n = 2
        gear = _floats(a.get("gear"), 1, [1])[0]
        A["act_trnnum"].append(n); A["act_trndof"].append(dofs); A["act_trncoef"].append([c * gear for c in coefs])
        if a["type"] == "motor":
            gain, bias = 1.0, [0, 0, 0]
        else:
            gain = _floats(a.get("gainprm"), 1, [1])[0]
            bias = _floats(a.get("biasprm"), 3, [0, 0, 0]) if a.get("biastype", "none") == "affine" else [0, 0, 0]
        A["act_gain"].append(gain); A["act_bias"].append(bias)
        cr, fr = a.get("ctrlrange"), a.get("forcerange")
        A["act_ctrllimited"].append(int(cr is not None)); A["act_ctrlrange"].append(_floats(cr, 2, [0, 0]))
        A["act_forcelimited"].append(int(fr is not None)); A["act_forcerange"].append(_floats(fr, 2, [0, 0]))
    counts["nu"] = len(model.actuators); counts["neq"] = len(eqs); counts["npair"] = len(pairs)
    # props
    nprop = getattr(model, "nprop", 0)
    A["body_propid"] = [-1] * counts["nbody"]; A["geom_propid"] = [-1] * counts["ngeom"]
    for p in range(nprop):
        g = names["geom"].index("prop_%d/prop_%d" % (p, p))
        b = A["geom_bodyid"][g]
        jb = b
        while A["body_jntnum"][jb] == 0:
            jb = A["body_parentid"][jb]
        j = A["body_jntadr"][jb]
        A["prop_bodyid"].append(b); A["prop_geomid"].append(g)
        A["prop_dofadr"].append(A["jnt_dofadr"][j]); A["prop_qposadr"].append(A["jnt_qposadr"][j])
        A["prop_mass"].append(A["body_mass"][b])
        A["body_propid"][b] = p; A["geom_propid"][g] = p
    counts["nprop"] = nprop
    first_prop_dof = A["prop_dofadr"][0] if nprop else counts["nv"]
    counts["narm"] = 7; counts["ngrip"] = first_prop_dof - 7
    counts["site_attachment"] = names["site"].index("panda nohand/attachment_site")
    pinch = "panda nohand/robotiq_2f85/pinch"
    counts["site_pinch"] = names["site"].index(pinch) if pinch in names["site"] else counts["site_attachment"]   # no gripper: the arm's own site
    counts["geom_table"] = names["geom"].index("table/table"); counts["geom_plane"] = names["geom"].index("ground")
    o = model.option
    grav = _floats(o.get("gravity"), 3, [0, 0, -9.81])
    integ = {"Euler": B.INT_EULER, "implicitfast": B.INT_IMPLICITFAST}[o.get("integrator", "Euler")]
    cone = {"pyramidal": B.CONE_PYRAMIDAL, "elliptic": B.CONE_ELLIPTIC}[o.get("cone", "pyramidal")]
    opts = dict(timestep=float(o.get("timestep", 0.002)), gravity_x=grav[0], gravity_y=grav[1], gravity_z=grav[2],
                impratio=float(o.get("impratio", 1)), tolerance=float(o.get("tolerance", 1e-8)),
                ls_tolerance=float(o.get("ls_tolerance", 0.01)), iterations=int(o.get("iterations", 100)),
                ls_iterations=int(o.get("ls_iterations", 50)), integrator=integ, cone=cone,
                ccd_iterations=int(o.get("ccd_iterations", 50)), ccd_tolerance=float(o.get("ccd_tolerance", 1e-6)),
                meaninertia=meaninertia)
    return counts, opts, A, names


def compile_rearrangement(n_props=4, prop_half_size=0.0155, fuse=True, **scene_kw):
    """Canonical padded scene: ``n_props`` cube props (per-env sizes override the blob's)."""
    sizes = [[prop_half_size] * 3 for _ in range(n_props)]
    model = build_scene(sizes, **scene_kw)
    counts, opts, A, names = compile_model(model, fuse=fuse)
    return B.pack(counts, opts, A), names


def compile_lasa_draw(fuse=True, **scene_kw):
    """The drawing scene of tasks/lasa_draw.py: 7 arm dofs, no gripper, no props, one cylinder tool piece, pyramidal cone."""
    model = build_draw_scene(**scene_kw)
    counts, opts, A, names = compile_model(model, fuse=fuse)
    return B.pack(counts, opts, A), names
