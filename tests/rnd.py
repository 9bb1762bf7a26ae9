"""ctypes wrapper of the CPU render oracle (oracle/b2p_render_oracle.c) -- checker only."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_L = None


def lib():
    global _L
    if _L is None:
        so = os.path.join(ROOT, "oracle", "libb2p_render_oracle.so")
        src = os.path.join(ROOT, "oracle", "b2p_render_oracle.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL)
        _L = C.CDLL(so)
        _L.b2p_render_oracle.argtypes = [C.c_char_p, C.c_size_t] + [C.c_void_p] * 8 + [C.c_int] * 4
    return _L


def render_blob(tag="rearrangement_n4"):
    with open(os.path.join(ROOT, "mujoco_robot_environments_b200", "assets", tag + ".b2pr"), "rb") as f:
        return f.read()


def header(blob):
    hd = np.frombuffer(blob[:64], np.int32)
    return dict(ngeom=int(hd[2]), nplane=int(hd[3]), nzone=int(hd[4]), W=int(hd[5]), H=int(hd[6]), nprop=int(hd[7]))


def render(blob, geom_xpos, geom_xmat, prop_rgba=None, prop_size=None, prop_active=None, window=None, out=None):
    """-> (rgb u8 [H,W,3], depth f32 [H,W], seg i32 [H,W]); ``window`` = (x0, y0, x1, y1) renders only that rectangle."""
    h = header(blob)
    xp = np.ascontiguousarray(geom_xpos, np.float64); xm = np.ascontiguousarray(geom_xmat, np.float64)
    rgba = None if prop_rgba is None else np.ascontiguousarray(prop_rgba, np.float32)
    ps = None if prop_size is None else np.ascontiguousarray(prop_size, np.float64)
    pa = None if prop_active is None else np.ascontiguousarray(prop_active, np.uint8)
    if out is None:
        out = (np.zeros((h["H"], h["W"], 3), np.uint8), np.zeros((h["H"], h["W"]), np.float32), np.full((h["H"], h["W"]), -2, np.int32))
    x0, y0, x1, y1 = window if window is not None else (0, 0, 0, 0)
    p = lambda a: None if a is None else a.ctypes.data
    rc = lib().b2p_render_oracle(blob, len(blob), p(xp), p(xm), p(rgba), p(ps), p(pa), p(out[0]), p(out[1]), p(out[2]), x0, y0, x1, y1)
    assert rc == 0
    return out
