"""Pin the oracle against real MuJoCo -- runs ONLY where `mujoco==3.2.7` (and the menagerie XML/meshes) are available.

MuJoCo is not installable in the build image or on the GPU box, so this script cannot run there;
it exists so that anyone with `pip install mujoco==3.2.7` can produce `tests/golden/mujoco_step_golden.npz` in the
schema of tests/golden/make_step_golden.py and turn "parity unpinned" into a pinned comparison:

  python tools/dump_mujoco_golden.py --menagerie /path/to/mujoco_menagerie

It (1) re-assembles the scene exactly as tools/compile_scene.py does (same MJCF writer), (2) loads it with
mujoco.MjModel.from_xml_string, (3) prints the model-constant diff against the committed blob (body_mass, body_inertia,
dof_invweight0, body_invweight0, geom_rbound, opt.*: the SURVEY 7.1-0 checklist), and (4) replays the seeded states /
controls of step_golden.npz with mj_step, storing qpos/qvel/qacc/contacts per step.
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    try:
        import mujoco
    except ImportError:
        sys.exit("mujoco is not installed here: this tool only runs where `pip install mujoco==3.2.7` is possible")
    from mujoco_robot_environments_b200.compile import blob as B
    from mujoco_robot_environments_b200.compile import scene
    import scenarios as sc

    ap = argparse.ArgumentParser()
    ap.add_argument("--menagerie", default=scene.DEFAULT_MENAGERIE)
    args = ap.parse_args()
    g = dict(np.load(os.path.join(ROOT, "tests", "golden", "step_golden.npz")))
    S, K = g["qpos"].shape[:2]
    out = {k: g[k] for k in ("state0", "prop_size", "prop_active", "ctrl")}
    for k in ("qpos", "qvel", "qacc"):
        out[k] = np.zeros_like(g[k])
    out["ncon"] = np.zeros((S, K), np.int32); out["contact_geom"] = -np.ones((S, K, 48, 2), np.int32); out["contact_dist"] = np.zeros((S, K, 48))
    for e in range(S):
        model = scene.build_scene([tuple(h) for h in g["prop_size"][e]], menagerie=args.menagerie)
        xml = model.to_xml_string() if hasattr(model, "to_xml_string") else None
        if xml is None:
            sys.exit("compile/mjcf.py Model has no MJCF writer in this checkout; add to_xml_string() before using this tool")
        m = mujoco.MjModel.from_xml_string(xml, getattr(model, "assets", None))
        if e == 0:
            counts, opts, arrays = B.unpack(sc.default_blob("rearrangement_n4_unfused"))
            print("opt: integrator %d cone %d impratio %g solver %d iterations %d tolerance %g" % (m.opt.integrator, m.opt.cone, m.opt.impratio, m.opt.solver, m.opt.iterations, m.opt.tolerance))
            for name, ref in (("body_mass", m.body_mass), ("body_inertia", m.body_inertia), ("dof_invweight0", m.dof_invweight0),
                              ("body_invweight0", m.body_invweight0), ("geom_rbound", m.geom_rbound)):
                if name in arrays and np.size(arrays[name]) == np.size(ref):
                    print("  %-16s max |blob - mujoco| = %.3e" % (name, np.abs(np.asarray(arrays[name]).reshape(np.shape(ref)) - ref).max()))
        d = mujoco.MjData(m)
        d.qpos[:] = g["state0"][e, :m.nq]; d.qvel[:] = g["state0"][e, sc.REC_QVEL:sc.REC_QVEL + m.nv]
        for p in range(4):
            if not g["prop_active"][e, p]:
                gid = mujoco.mj_name2id(m, mujoco.mjtObj.mjOBJ_GEOM, "prop_%d/prop_%d" % (p, p))
                m.geom_contype[gid] = 0; m.geom_conaffinity[gid] = 0
        d.ctrl[:] = g["ctrl"][e]
        for k in range(K):
            mujoco.mj_step(m, d)
            out["qpos"][e, k] = d.qpos; out["qvel"][e, k] = d.qvel; out["qacc"][e, k] = d.qacc
            act = [c for c in d.contact[:d.ncon] if c.exclude == 0]
            out["ncon"][e, k] = len(act)
            for i, c in enumerate(act[:48]):
                out["contact_geom"][e, k, i] = (c.geom1, c.geom2); out["contact_dist"][e, k, i] = c.dist
        print("scene %d: max |qpos_mujoco - qpos_oracle| after %d steps = %.3e" % (e, K, np.abs(out["qpos"][e, -1] - g["qpos"][e, -1]).max()))
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "mujoco_step_golden.npz"), **out)


if __name__ == "__main__":
    main()
