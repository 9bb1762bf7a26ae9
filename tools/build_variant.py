"""Build a variant of libb2p.so with extra nvcc flags into mujoco_robot_environments_b200/csrc/variants/libb2p_<name>.so (developer aid for
A/B runs on the GPU box: `B2P_LIB=<path> python bench.py ...` loads it instead of the product build; never a fallback).

  python tools/build_variant.py prof -DB2P_PROF
  python tools/build_variant.py w13 -DB2P_WARPS_F32=13 -DB2P_JPS_F32=384
"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mujoco_robot_environments_b200 import ffi

name, extra = sys.argv[1], sys.argv[2:]
vdir = os.path.join(ffi.CSRC, "variants"); bdir = os.path.join(ffi.CSRC, "build", name)
os.makedirs(vdir, exist_ok=True); os.makedirs(bdir, exist_ok=True)
procs, objs = [], []
for u in ffi.UNITS:
    obj = os.path.join(bdir, u.replace(".cu", ".o")); objs.append(obj)
    cmd = ["nvcc"] + ffi.NVCC_FLAGS + ffi.UNIT_FLAGS.get(u, []) + extra + ["-c", "-o", obj, os.path.join(ffi.CSRC, u)]
    procs.append((cmd, subprocess.Popen(cmd)))
for cmd, p in procs:
    if p.wait() != 0:
        raise SystemExit("failed: " + " ".join(cmd))
out = os.path.join(vdir, "libb2p_%s.so" % name)
subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", out] + objs)
print(out)
