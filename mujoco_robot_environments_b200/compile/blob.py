"""Flat, versioned model blob ("B2PM") shared by the CPU oracle and the CUDA kernels.

The blob mirrors the subset of MuJoCo's ``mjModel`` that the rearrangement model
class needs: hinge + free joints, box / plane / convex-mesh
geoms, ``connect`` + ``joint`` equalities, one fixed tendon folded into the
actuator transmission, ``motor`` + affine ``general`` actuators.

Layout (little endian):
  int32  header[64]     magic, version, counts (see COUNTS)
  float64 opt[32]       options (see OPTS)
  int32  table[2*NF]    (offset_in_bytes, n_elements) per field in FIELDS order
  ...                   8-byte aligned field payloads (int32 or float64)

``include/b2p_model_layout.h`` is generated from this file by
``tools/gen_layout_header.py`` so that C / CUDA and Python can never disagree.
"""
from __future__ import annotations

import numpy as np

MAGIC = 0x4D503242  # "B2PM"
VERSION = 3

COUNTS = [
    "magic", "version", "nq", "nv", "nu", "nbody", "njnt", "ngeom", "nsite", "neq",
    "npair", "nmesh", "nmeshvert", "nprop", "narm", "ngrip", "nexclude_unused", "ntree_unused",
    "site_attachment", "site_pinch", "geom_table", "geom_plane", "nmeshnbr",
]

OPTS = [
    "timestep", "gravity_x", "gravity_y", "gravity_z", "impratio", "tolerance", "ls_tolerance",
    "iterations", "ls_iterations", "integrator", "cone", "ccd_iterations", "ccd_tolerance",
    "meaninertia",
]

I, F = "i", "f"
# (name, dtype, elements-per-item, count-key)
FIELDS = [
    # bodies
    ("body_parentid", I, 1, "nbody"), ("body_rootid", I, 1, "nbody"), ("body_weldid", I, 1, "nbody"),
    ("body_jntnum", I, 1, "nbody"), ("body_jntadr", I, 1, "nbody"),
    ("body_dofnum", I, 1, "nbody"), ("body_dofadr", I, 1, "nbody"),
    ("body_propid", I, 1, "nbody"),
    ("body_pos", F, 3, "nbody"), ("body_quat", F, 4, "nbody"),
    ("body_ipos", F, 3, "nbody"), ("body_iquat", F, 4, "nbody"),
    ("body_mass", F, 1, "nbody"), ("body_inertia", F, 3, "nbody"),
    ("body_invweight0", F, 2, "nbody"),
    # joints
    ("jnt_type", I, 1, "njnt"), ("jnt_qposadr", I, 1, "njnt"), ("jnt_dofadr", I, 1, "njnt"),
    ("jnt_bodyid", I, 1, "njnt"), ("jnt_limited", I, 1, "njnt"),
    ("jnt_pos", F, 3, "njnt"), ("jnt_axis", F, 3, "njnt"), ("jnt_stiffness", F, 1, "njnt"),
    ("jnt_range", F, 2, "njnt"), ("jnt_margin", F, 1, "njnt"),
    ("jnt_solref", F, 2, "njnt"), ("jnt_solimp", F, 5, "njnt"),
    # dofs
    ("dof_bodyid", I, 1, "nv"), ("dof_jntid", I, 1, "nv"), ("dof_parentid", I, 1, "nv"),
    ("dof_armature", F, 1, "nv"), ("dof_damping", F, 1, "nv"), ("dof_invweight0", F, 1, "nv"),
    ("qpos0", F, 1, "nq"), ("qpos_spring", F, 1, "nq"),
    # geoms (collidable only; visual geoms are stripped by the compiler)
    ("geom_type", I, 1, "ngeom"), ("geom_bodyid", I, 1, "ngeom"), ("geom_condim", I, 1, "ngeom"),
    ("geom_priority", I, 1, "ngeom"), ("geom_meshid", I, 1, "ngeom"), ("geom_propid", I, 1, "ngeom"),
    ("geom_size", F, 3, "ngeom"), ("geom_pos", F, 3, "ngeom"), ("geom_quat", F, 4, "ngeom"),
    ("geom_friction", F, 3, "ngeom"), ("geom_solref", F, 2, "ngeom"), ("geom_solimp", F, 5, "ngeom"),
    ("geom_solmix", F, 1, "ngeom"), ("geom_margin", F, 1, "ngeom"), ("geom_gap", F, 1, "ngeom"),
    ("geom_rbound", F, 1, "ngeom"), ("geom_aabb", F, 6, "ngeom"),
    ("geom_invweight0", F, 2, "ngeom"),
    # convex meshes (hull vertices in the geom frame, vertex adjacency for hill climbing)
    ("mesh_vertadr", I, 1, "nmesh"), ("mesh_vertnum", I, 1, "nmesh"),
    ("mesh_vert", F, 3, "nmeshvert"),
    ("mesh_nbradr", I, 1, "nmeshvert"), ("mesh_nbrnum", I, 1, "nmeshvert"),
    ("mesh_nbr", I, 1, "nmeshnbr"),
    # sites
    ("site_bodyid", I, 1, "nsite"), ("site_pos", F, 3, "nsite"), ("site_quat", F, 4, "nsite"),
    # candidate geom pairs (after static filters: contype/conaffinity, weld, parent, exclude)
    ("pair_geom1", I, 1, "npair"), ("pair_geom2", I, 1, "npair"),
    # equality constraints
    ("eq_type", I, 1, "neq"), ("eq_obj1id", I, 1, "neq"), ("eq_obj2id", I, 1, "neq"),
    ("eq_data", F, 11, "neq"), ("eq_solref", F, 2, "neq"), ("eq_solimp", F, 5, "neq"),
    # actuators (fixed tendon folded into a two-dof transmission)
    ("act_trnnum", I, 1, "nu"), ("act_trndof", I, 2, "nu"),
    ("act_ctrllimited", I, 1, "nu"), ("act_forcelimited", I, 1, "nu"),
    ("act_trncoef", F, 2, "nu"), ("act_gain", F, 1, "nu"), ("act_bias", F, 3, "nu"),
    ("act_ctrlrange", F, 2, "nu"), ("act_forcerange", F, 2, "nu"),
    # props (free bodies whose size is per-env data)
    ("prop_bodyid", I, 1, "nprop"), ("prop_geomid", I, 1, "nprop"),
    ("prop_dofadr", I, 1, "nprop"), ("prop_qposadr", I, 1, "nprop"),
    ("prop_mass", F, 1, "nprop"),
]

# enums shared with C
JNT_FREE, JNT_HINGE = 0, 3
GEOM_PLANE, GEOM_SPHERE, GEOM_CYLINDER, GEOM_BOX, GEOM_MESH = 0, 2, 5, 6, 7
EQ_CONNECT, EQ_JOINT = 0, 2
INT_EULER, INT_IMPLICITFAST = 0, 3
CONE_PYRAMIDAL, CONE_ELLIPTIC = 0, 1


def pack(counts: dict, opts: dict, arrays: dict) -> bytes:
    """Serialise a compiled model into the B2PM byte layout."""
    header = np.zeros(64, np.int32)
    counts = dict(counts)
    counts["magic"], counts["version"] = MAGIC, VERSION
    for i, k in enumerate(COUNTS):
        header[i] = int(counts.get(k, 0))
    opt = np.zeros(32, np.float64)
    for i, k in enumerate(OPTS):
        opt[i] = float(opts[k])
    nf = len(FIELDS)
    table = np.zeros(2 * nf, np.int32)
    payload = []
    off = header.nbytes + opt.nbytes + table.nbytes
    assert off % 8 == 0
    for fi, (name, dt, per, ckey) in enumerate(FIELDS):
        n = int(counts.get(ckey, 0)) * per
        a = np.asarray(arrays[name], dtype=np.int32 if dt == I else np.float64).reshape(-1)
        if a.size != n:
            raise ValueError(f"field {name}: expected {n} elements, got {a.size}")
        table[2 * fi], table[2 * fi + 1] = off, n
        b = a.tobytes()
        pad = (-len(b)) % 8
        payload.append(b + b"\0" * pad)
        off += len(b) + pad
    return header.tobytes() + opt.tobytes() + table.tobytes() + b"".join(payload)


def unpack(buf: bytes):
    """Inverse of :func:`pack` -> (counts, opts, arrays)."""
    header = np.frombuffer(buf, np.int32, 64, 0)
    if header[0] != MAGIC or header[1] != VERSION:
        raise ValueError("not a B2PM v%d blob" % VERSION)
    counts = {k: int(header[i]) for i, k in enumerate(COUNTS)}
    opt = np.frombuffer(buf, np.float64, 32, 256)
    opts = {k: float(opt[i]) for i, k in enumerate(OPTS)}
    nf = len(FIELDS)
    table = np.frombuffer(buf, np.int32, 2 * nf, 256 + 256)
    arrays = {}
    for fi, (name, dt, per, ckey) in enumerate(FIELDS):
        off, n = int(table[2 * fi]), int(table[2 * fi + 1])
        a = np.frombuffer(buf, np.int32 if dt == I else np.float64, n, off).copy()
        arrays[name] = a.reshape(-1, per) if per > 1 else a
    return counts, opts, arrays


def c_header() -> str:
    """Text of include/b2p_model_layout.h (generated; do not edit by hand)."""
    out = ["/* GENERATED by tools/gen_layout_header.py from",
           " * mujoco_robot_environments_b200/compile/blob.py -- do not edit. */",
           "#ifndef B2P_MODEL_LAYOUT_H", "#define B2P_MODEL_LAYOUT_H",
           f"#define B2PM_MAGIC 0x{MAGIC:08X}", f"#define B2PM_VERSION {VERSION}", ""]
    out.append("enum b2pm_count {")
    out += [f"  B2PM_C_{k.upper()} = {i}," for i, k in enumerate(COUNTS)]
    out.append("};\nenum b2pm_opt {")
    out += [f"  B2PM_O_{k.upper()} = {i}," for i, k in enumerate(OPTS)]
    out.append("};\nenum b2pm_field {")
    out += [f"  B2PM_F_{n.upper()} = {i},  /* {'int32' if d == I else 'float64'} x{p} per {c} */"
            for i, (n, d, p, c) in enumerate(FIELDS)]
    out.append(f"  B2PM_NFIELD = {len(FIELDS)}\n}};")
    out.append("#define B2PM_HEADER_BYTES 256\n#define B2PM_OPT_BYTES 256")
    out.append("#define B2PM_TABLE_OFFSET 512")
    out.append("enum { B2P_JNT_FREE = 0, B2P_JNT_HINGE = 3 };")
    out.append("enum { B2P_GEOM_PLANE = 0, B2P_GEOM_SPHERE = 2, B2P_GEOM_CYLINDER = 5, B2P_GEOM_BOX = 6, B2P_GEOM_MESH = 7 };")
    out.append("enum { B2P_EQ_CONNECT = 0, B2P_EQ_JOINT = 2 };")
    out.append("enum { B2P_INT_EULER = 0, B2P_INT_IMPLICITFAST = 3 };")
    out.append("enum { B2P_CONE_PYRAMIDAL = 0, B2P_CONE_ELLIPTIC = 1 };")
    out.append("#endif")
    return "\n".join(out) + "\n"
