"""Minimal MJCF reader for the rearrangement model class.

Reads what ``RearrangementEnv.__init__`` feeds to the MuJoCo compiler
(reference: mujoco_robot_environments/tasks/rearrangement.py:54-210): default
classes / childclass resolution, bodies, inertials, hinge + free joints, geoms,
sites, ``<contact><exclude>``, one ``<tendon><fixed>``, ``<equality>``
(``connect`` / ``joint``) and ``<actuator>`` (``general`` / ``motor``).
It is NOT a general MJCF implementation.

dm_control's ``attach`` semantics (name scoping ``<model>/<element>``, frame body
at the attachment site, merged ``<option>``) are reproduced by
:func:`Model.attach`.
"""
from __future__ import annotations

import os
import struct
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np


def _floats(s, n=None, default=None):
    if s is None:
        return None if default is None else np.array(default, float)
    v = np.array([float(x) for x in str(s).split()], float)
    if n is not None and v.size < n and default is not None:
        d = np.array(default, float)
        d[: v.size] = v
        v = d
    return v


# ---------------------------------------------------------------- quaternion helpers (w,x,y,z)
def quat_mul(a, b):
    return np.array([
        a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3],
        a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
        a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1],
        a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]])


def quat_to_mat(q):
    w, x, y, z = q
    return np.array([
        [w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y)],
        [2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x)],
        [2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z]])


def mat_to_quat(m):
    """Branch order of MuJoCo's mju_mat2Quat."""
    m = np.asarray(m, float).reshape(3, 3)
    q = np.zeros(4)
    if m[0, 0] + m[1, 1] + m[2, 2] > 0:
        q[0] = 0.5 * np.sqrt(1 + m[0, 0] + m[1, 1] + m[2, 2])
        q[1] = 0.25 * (m[2, 1] - m[1, 2]) / q[0]
        q[2] = 0.25 * (m[0, 2] - m[2, 0]) / q[0]
        q[3] = 0.25 * (m[1, 0] - m[0, 1]) / q[0]
    elif m[0, 0] > m[1, 1] and m[0, 0] > m[2, 2]:
        q[1] = 0.5 * np.sqrt(1 + m[0, 0] - m[1, 1] - m[2, 2])
        q[0] = 0.25 * (m[2, 1] - m[1, 2]) / q[1]
        q[2] = 0.25 * (m[0, 1] + m[1, 0]) / q[1]
        q[3] = 0.25 * (m[0, 2] + m[2, 0]) / q[1]
    elif m[1, 1] > m[2, 2]:
        q[2] = 0.5 * np.sqrt(1 - m[0, 0] + m[1, 1] - m[2, 2])
        q[0] = 0.25 * (m[0, 2] - m[2, 0]) / q[2]
        q[1] = 0.25 * (m[0, 1] + m[1, 0]) / q[2]
        q[3] = 0.25 * (m[1, 2] + m[2, 1]) / q[2]
    else:
        q[3] = 0.5 * np.sqrt(1 - m[0, 0] - m[1, 1] + m[2, 2])
        q[0] = 0.25 * (m[1, 0] - m[0, 1]) / q[3]
        q[1] = 0.25 * (m[0, 2] + m[2, 0]) / q[3]
        q[2] = 0.25 * (m[1, 2] + m[2, 1]) / q[3]
    return q / np.linalg.norm(q)


def quat_normalize(q):
    q = np.asarray(q, float)
    return q / np.linalg.norm(q)


# ---------------------------------------------------------------- meshes
def load_mesh(path: str):
    """-> (vertices [n,3], faces [m,3]) from binary STL or OBJ."""
    if path.lower().endswith(".stl"):
        b = open(path, "rb").read()
        n = struct.unpack("<I", b[80:84])[0]
        if 84 + 50 * n != len(b):
            raise ValueError("only binary STL supported: " + path)
        rec = np.frombuffer(b[84:], dtype=np.dtype([("n", "<f4", 3), ("v", "<f4", 9), ("a", "<u2")]))
        v = rec["v"].reshape(-1, 3).astype(np.float64)
        f = np.arange(3 * n).reshape(-1, 3)
        return v, f
    verts, faces = [], []
    for line in open(path):
        p = line.split()
        if not p:
            continue
        if p[0] == "v":
            verts.append([float(x) for x in p[1:4]])
        elif p[0] == "f":
            idx = [int(t.split("/")[0]) - 1 for t in p[1:]]
            for k in range(1, len(idx) - 1):
                faces.append([idx[0], idx[k], idx[k + 1]])
    return np.array(verts, float), np.array(faces, int)


def mesh_mass_properties(v, f):
    """Signed-tetrahedra volume integrals of a closed triangle mesh.

    -> (volume, com[3], inertia 3x3 about com for unit density)."""
    a, b, c = v[f[:, 0]], v[f[:, 1]], v[f[:, 2]]
    det = np.einsum("ij,ij->i", a, np.cross(b, c))
    vol = det.sum() / 6.0
    if vol < 0:  # inward-facing winding
        det, vol = -det, -vol
    com = ((a + b + c) * det[:, None]).sum(0) / (24.0 * vol)
    # second moments: integral of x_i x_j over each tetra (origin, a, b, c)
    S = np.zeros((3, 3))
    for p, q in ((a, a), (b, b), (c, c)):
        S += np.einsum("n,ni,nj->ij", det, p, q) * 2
    for p, q in ((a, b), (a, c), (b, c)):
        S += np.einsum("n,ni,nj->ij", det, p, q) + np.einsum("n,ni,nj->ij", det, q, p)
    S /= 120.0
    S -= vol * np.outer(com, com)  # shift second moments to com
    inertia = np.trace(S) * np.eye(3) - S
    return vol, com, inertia


def convex_hull(v):
    """-> (hull vertices [k,3], neighbour lists) using qhull (what MuJoCo's compiler uses)."""
    from scipy.spatial import ConvexHull
    h = ConvexHull(v)
    idx = np.sort(h.vertices)
    remap = -np.ones(len(v), int)
    remap[idx] = np.arange(len(idx))
    nbr = [set() for _ in idx]
    for s in h.simplices:
        s = remap[s]
        for i in range(3):
            for j in range(3):
                if i != j:
                    nbr[s[i]].add(int(s[j]))
    return v[idx].copy(), [sorted(x) for x in nbr]


# ---------------------------------------------------------------- element tree with defaults
@dataclass
class Body:
    name: str
    pos: np.ndarray
    quat: np.ndarray
    parent: Optional["Body"] = None
    children: List["Body"] = field(default_factory=list)
    joints: List[dict] = field(default_factory=list)
    geoms: List[dict] = field(default_factory=list)
    sites: List[dict] = field(default_factory=list)
    inertial: Optional[dict] = None


class Model:
    """A parsed MJCF model (or the composed arena)."""

    def __init__(self, name="arena"):
        self.name = name
        self.world = Body("world", np.zeros(3), np.array([1.0, 0, 0, 0]))
        self.option: Dict[str, object] = {}
        self.meshes: Dict[str, dict] = {}
        self.excludes: List[tuple] = []
        self.tendons: List[dict] = []
        self.equalities: List[dict] = []
        self.actuators: List[dict] = []

    # ---- parsing
    @classmethod
    def from_path(cls, path: str) -> "Model":
        root = ET.parse(path).getroot()
        m = cls(root.get("model", "model"))
        comp = root.find("compiler")
        meshdir = os.path.join(os.path.dirname(path), comp.get("meshdir", "") if comp is not None else "")
        if comp is not None and comp.get("angle", "degree") != "radian":
            raise ValueError("only angle=radian models are supported")
        m.autolimits = comp is not None and comp.get("autolimits", "false") == "true"
        opt = root.find("option")
        if opt is not None:
            m.option.update(opt.attrib)
        # defaults: class name -> {tag: attribs}, with inheritance
        m.defaults = {"main": {}}

        def walk_default(el, parent):
            cname = el.get("class", "main")
            d = {t: dict(a) for t, a in m.defaults.get(parent, {}).items()} if parent else {}
            for ch in el:
                if ch.tag != "default":
                    d.setdefault(ch.tag, {})
                    d[ch.tag] = {**d[ch.tag], **ch.attrib}
            m.defaults[cname] = d
            for ch in el:
                if ch.tag == "default":
                    walk_default(ch, cname)
        for d in root.findall("default"):
            walk_default(d, None)

        def resolve(el, tag, childclass):
            c = el.get("class", childclass or "main")
            a = dict(m.defaults.get(c, {}).get(tag, {}))
            a.update({k: v for k, v in el.attrib.items() if k != "class"})
            return a

        for asset in root.findall("asset"):
            for me in asset.findall("mesh"):
                a = resolve(me, "mesh", None)
                name = a.get("name") or os.path.splitext(os.path.basename(a["file"]))[0]
                m.meshes[name] = {"file": os.path.join(meshdir, a["file"]),
                                  "scale": _floats(a.get("scale"), 3, [1, 1, 1])}

        def parse_body(el, parent: Body, childclass):
            for ch in el:
                if ch.tag == "body":
                    cc = ch.get("childclass", childclass)
                    b = Body(ch.get("name", ""), _floats(ch.get("pos"), 3, [0, 0, 0]),
                             quat_normalize(_floats(ch.get("quat"), 4, [1, 0, 0, 0])), parent)
                    parent.children.append(b)
                    parse_body(ch, b, cc)
                elif ch.tag == "inertial":
                    parent.inertial = dict(ch.attrib)
                elif ch.tag in ("joint", "freejoint"):
                    a = resolve(ch, "joint", childclass) if ch.tag == "joint" else dict(ch.attrib, type="free")
                    parent.joints.append(a)
                elif ch.tag == "geom":
                    parent.geoms.append(resolve(ch, "geom", childclass))
                elif ch.tag == "site":
                    parent.sites.append(resolve(ch, "site", childclass))
        wb = root.find("worldbody")
        if wb is not None:
            parse_body(wb, m.world, None)
        con = root.find("contact")
        if con is not None:
            for e in con.findall("exclude"):
                m.excludes.append((e.get("body1"), e.get("body2")))
        ten = root.find("tendon")
        if ten is not None:
            for t in ten.findall("fixed"):
                m.tendons.append({"name": t.get("name"),
                                  "joints": [(j.get("joint"), float(j.get("coef"))) for j in t.findall("joint")]})
        eq = root.find("equality")
        if eq is not None:
            for e in eq:
                m.equalities.append(dict(resolve(e, e.tag, None), type=e.tag))
        act = root.find("actuator")
        if act is not None:
            for e in act:
                m.actuators.append(dict(resolve(e, e.tag, e.get("class")), type=e.tag))
        return m

    # ---- traversal
    def bodies(self) -> List[Body]:
        out = []

        def rec(b):
            out.append(b)
            for c in b.children:
                rec(c)
        rec(self.world)
        return out

    def find_body(self, name) -> Body:
        for b in self.bodies():
            if b.name == name:
                return b
        raise KeyError(name)

    def find_site(self, name):
        for b in self.bodies():
            for s in b.sites:
                if s.get("name") == name:
                    return b, s
        raise KeyError(name)

    def all_joints(self):
        return [j for b in self.bodies() for j in b.joints]

    # ---- dm_control-style composition
    def _scope(self, prefix: str):
        """Prefix every name with ``prefix`` (dm_control namescope)."""
        for b in self.bodies():
            if b is not self.world:
                b.name = prefix + b.name
            for coll in (b.joints, b.geoms, b.sites):
                for e in coll:
                    if e.get("name"):
                        e["name"] = prefix + e["name"]
            for g in b.geoms:
                if g.get("mesh"):
                    g["mesh"] = prefix + g["mesh"]
        self.meshes = {prefix + k: v for k, v in self.meshes.items()}
        self.excludes = [(prefix + a, prefix + b) for a, b in self.excludes]
        for t in self.tendons:
            t["name"] = prefix + t["name"]
            t["joints"] = [(prefix + j, c) for j, c in t["joints"]]
        for e in self.equalities:
            for k in ("body1", "body2", "joint1", "joint2"):
                if e.get(k):
                    e[k] = prefix + e[k]
        for a in self.actuators:
            for k in ("name", "joint", "tendon"):
                if a.get(k):
                    a[k] = prefix + a[k]

    def attach(self, child: "Model", site_name: Optional[str] = None, pos=None, free=False) -> Body:
        """dm_control ``Entity.attach`` / ``Arena.add_free_entity``: the child's worldbody
        becomes a frame body ``<child>/`` under the body that owns ``site_name``.
        ``<option>`` attributes are merged (conflicts raise, as in dm_control)."""
        prefix = child.name + "/"
        child._scope(prefix)
        if site_name is not None:
            parent, site = self.find_site(site_name)
            fpos = _floats(site.get("pos"), 3, [0, 0, 0])
            fquat = quat_normalize(_floats(site.get("quat"), 4, [1, 0, 0, 0]))
        else:
            parent, fpos, fquat = self.world, np.zeros(3) if pos is None else np.asarray(pos, float), np.array([1.0, 0, 0, 0])
        frame = Body(prefix, fpos, fquat, parent)
        if free:
            frame.joints.append({"type": "free", "name": prefix})
        frame.children = child.world.children
        frame.geoms, frame.sites = child.world.geoms, child.world.sites
        for c in frame.children:
            c.parent = frame
        parent.children.append(frame)
        for k, v in child.option.items():
            if k in self.option and self.option[k] != v:
                raise ValueError(f"conflicting <option {k}>: {self.option[k]} vs {v}")
            self.option[k] = v
        self.meshes.update(child.meshes)
        self.excludes += child.excludes
        self.tendons += child.tendons
        self.equalities += child.equalities
        self.actuators += child.actuators
        return frame
