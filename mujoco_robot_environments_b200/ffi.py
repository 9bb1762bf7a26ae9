"""ctypes loader for libb2p.so (the C-ABI of include/b2p.h) and its in-tree build.

There is no CPU fallback: if the CUDA extension is missing or fails to load, importing the
physics backend raises.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.environ.get("B2P_LIB") or os.path.join(CSRC, "libb2p.so")      # B2P_LIB: a debug build (tools/), never a fallback
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC"]
UNITS = ("b2p_api.cu", "b2p_kernel_f32.cu", "b2p_kernel_f64.cu", "b2p_render.cu")
# per-unit flags.  fp32 throughput kernel: approximate division / square root (<= 2 ulp) and flush-to-zero, but NOT -use_fast_math:
# its __sinf / __cosf / __powf substitutions (absolute error 2^-21) made the Newton solve of ill-conditioned grasps wander (39
# iterations and a wrong qacc on tests/golden/bad_state_fp32_0.json, against 2 iterations with these flags) for 3 % of throughput.
# The ray-caster is built without FMA contraction: every pixel is bit-identical to the plain-C restatement (oracle/b2p_render_oracle.c)
UNIT_FLAGS = {"b2p_kernel_f32.cu": ["--prec-div=false", "--prec-sqrt=false", "--ftz=true"], "b2p_render.cu": ["-fmad=false"]}

# enums of include/b2p.h
F64, F32 = 0, 1
Q = dict(NQ=0, NV=1, NU=2, NPROP=3, STATE_STRIDE=4, OUT_STRIDE=5, MAXPROP=6, REC_QPOS=7, REC_QVEL=8, REC_QACC_WARMSTART=9,
         REC_TIME=10, LAUNCHES=11, SMEM_BYTES=12, BLOCKS_PER_SM=13, THREADS=14, GRID=15, MAXCON=16, NGEOM=17, ORDER_LAUNCHES=18)
O = dict(SITE_XPOS=0, SITE_XMAT=1, QACC=2, QFRC_BIAS=3, CTRL=4, NCON=5, NEFC=6, NITER=7, CONTACT=8, MARM=9, GEOM_XPOS=10,
         GEOM_XMAT=11, QACC_SMOOTH=12, EFC_FORCE=13)
SYMBOLS = ["b2p_create", "b2p_destroy", "b2p_query", "b2p_out_offset", "b2p_bind_state", "b2p_bind_out", "b2p_set_env_params", "b2p_set_out_mask", "b2p_set_solver", "b2p_set_env_prop_types",
           "b2p_osc_config", "b2p_set_env_gains", "b2p_set_env_mask", "b2p_set_full_contacts", "b2p_step_physics", "b2p_step_osc", "b2p_settle", "b2p_forward", "b2p_place_search", "b2p_place_props", "b2p_get",
           "b2p_status", "b2p_status_set", "b2p_prof_read", "b2p_fma_peak", "b2p_last_error", "b2p_step_physics_noalloc",
           "b2p_render_create", "b2p_render_destroy", "b2p_render_query", "b2p_render"]          # include/b2p_render.h
GEOM = dict(box=6, sphere=2, cylinder=5)            # B2P_GEOM_* of include/b2p_model_layout.h (MuJoCo's mjtGeom numbering)
SOLVER = dict(newton=0, pgs=1)
OUT = dict(SITE=1, CONTACT=2, MARM=4, GEOM=8, BIAS=16, DYN=32, EFC=64, ALL=127)
RQ = dict(WIDTH=0, HEIGHT=1, NGEOM=2, LAUNCHES=3, TILES_COPIED=4, TILES_CAST=5)


def sources():
    return [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cu", ".h"))] + \
           [os.path.join(_HERE, "..", "include", f) for f in ("b2p.h", "b2p_model_layout.h", "b2p_render.h")]


def build(force: bool = False, verbose: bool = False, extra=()) -> str:
    """Compile libb2p.so in-tree for sm_100a (nvcc cross-compiles without a GPU).  The translation units are
    compiled in parallel (the fp32 and fp64 kernel instantiations dominate) and linked into one shared library."""
    if not force and os.path.exists(LIB_PATH) and os.path.getmtime(LIB_PATH) >= max(os.path.getmtime(s) for s in sources()):
        return LIB_PATH
    bdir = os.path.join(CSRC, "build")
    os.makedirs(bdir, exist_ok=True)
    procs, objs = [], []
    for u in UNITS:
        obj = os.path.join(bdir, u.replace(".cu", ".o"))
        objs.append(obj)
        cmd = ["nvcc"] + NVCC_FLAGS + UNIT_FLAGS.get(u, []) + list(extra) + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj, os.path.join(CSRC, u)]
        procs.append((cmd, subprocess.Popen(cmd)))
    for cmd, p in procs:
        if p.wait() != 0:
            raise subprocess.CalledProcessError(p.returncode, cmd)
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB_PATH] + objs)
    return LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback for the physics backend)")
    L = C.CDLL(LIB_PATH)
    vp, i32, u8p = C.c_void_p, C.c_int, C.c_void_p
    L.b2p_create.argtypes = [C.c_char_p, C.c_size_t, i32, i32, i32, C.POINTER(vp)]
    L.b2p_destroy.argtypes = [vp]
    L.b2p_query.argtypes = [vp, i32]
    L.b2p_out_offset.argtypes = [i32, C.POINTER(i32), C.POINTER(i32)]
    L.b2p_bind_state.argtypes = [vp, vp]
    L.b2p_bind_out.argtypes = [vp, vp]
    L.b2p_set_env_params.argtypes = [vp, vp, u8p]
    L.b2p_set_out_mask.argtypes = [vp, i32]
    L.b2p_set_solver.argtypes = [vp, i32]
    L.b2p_set_env_prop_types.argtypes = [vp, u8p]
    L.b2p_osc_config.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_double, C.c_double]
    L.b2p_set_full_contacts.argtypes = [vp, i32]
    L.b2p_set_env_gains.argtypes = [vp, vp]
    L.b2p_set_env_mask.argtypes = [vp, vp]
    L.b2p_step_physics.argtypes = [vp, vp, i32, vp]
    L.b2p_step_physics_noalloc.argtypes = [vp, vp, vp, i32, vp]
    L.b2p_step_osc.argtypes = [vp, vp, u8p, i32, i32, u8p, vp]
    L.b2p_settle.argtypes = [vp, u8p, i32, i32, C.c_double, C.c_double, vp, vp]
    L.b2p_forward.argtypes = [vp, i32, vp]
    L.b2p_place_search.argtypes = [vp, vp, vp, u8p, i32, C.c_double, i32, C.c_uint, vp, vp]
    L.b2p_place_props.argtypes = [vp, vp, vp, u8p, i32, C.c_double, i32, C.c_uint, vp, vp, vp]
    L.b2p_get.argtypes = [vp, i32, vp, vp]
    L.b2p_status.argtypes = [vp, vp, i32, vp]
    L.b2p_status_set.argtypes = [vp, vp, vp]
    L.b2p_prof_read.argtypes = [vp, C.POINTER(C.c_ulonglong), i32]
    L.b2p_fma_peak.argtypes = [i32, i32, C.POINTER(C.c_double)]
    L.b2p_last_error.restype = C.c_char_p
    L.b2p_render_create.argtypes = [C.c_char_p, C.c_size_t, i32, C.POINTER(vp)]
    L.b2p_render_destroy.argtypes = [vp]
    L.b2p_render_query.argtypes = [vp, i32]
    L.b2p_render.argtypes = [vp, i32, vp, vp, i32, i32, vp, vp, vp, vp, vp, vp, vp]
    for s in SYMBOLS:
        getattr(L, s)  # raises AttributeError if the library does not export what include/b2p.h declares
    _lib = L
    return L


def check(rc: int, what: str = "b2p"):
    if rc != 0:
        raise RuntimeError(f"{what} failed ({rc}): {lib().b2p_last_error().decode()}")
