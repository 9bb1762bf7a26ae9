"""ctypes wrapper around the host build of the kernel templates (tests/hostemu). Test infrastructure only."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MODE_PHYSICS, MODE_OSC, MODE_SETTLE, MODE_FORWARD = 0, 1, 2, 3
_lib = None


class EmuArgs(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("mode", "n_env", "n_steps", "control_steps", "state_stride", "full_contacts", "settle_min_steps", "out_stride")] + \
               [("settle_qvel_tol", C.c_double), ("settle_qacc_tol", C.c_double)] + \
               [(n, C.c_void_p) for n in ("state", "ctrl", "target", "grip", "prop_size", "prop_active", "env_mask", "converged", "status", "steps_done", "out", "gains")] + \
               [("solver", C.c_int), ("prop_type", C.c_void_p)]


def lib():
    global _lib
    if _lib is None:
        so = os.path.join(ROOT, "tests", "_hostemu", "libb2p_hostemu.so")
        csrc = os.path.join(ROOT, "mujoco_robot_environments_b200", "csrc")
        srcs = [os.path.join(ROOT, "tests", "hostemu", "b2p_hostemu.cpp")] + [os.path.join(csrc, f) for f in sorted(os.listdir(csrc)) if f.endswith(".h")]
        if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(s) for s in srcs):
            os.makedirs(os.path.dirname(so), exist_ok=True)
            subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-o", so, srcs[0]])
        L = C.CDLL(so)
        L.emu_create.restype = C.c_void_p
        L.emu_create.argtypes = [C.c_char_p, C.c_size_t, C.c_int]
        L.emu_run.argtypes = [C.c_void_p, C.c_int, C.POINTER(EmuArgs)]
        L.emu_destroy.argtypes = [C.c_void_p, C.c_int]
        L.emu_osc_config.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_double, C.c_double]
        _lib = L
    return _lib


def out_slices():
    """Field -> slice of the derived-output record (csrc/b2p_core.h OutLayout)."""
    MAXNV, MAXU, MAXCON, MAXTD, MAXG, MAXEFC = 40, 8, 48, 16, 32, 160
    o, d = 0, {}
    for name, n in (("site_xpos", 6), ("site_xmat", 18), ("qacc", MAXNV), ("qfrc_bias", MAXNV), ("ctrl", MAXU), ("ncon", 1), ("nefc", 1),
                    ("niter", 1), ("contact", MAXCON * 10), ("Marm", MAXTD * MAXTD), ("gpos", MAXG * 3), ("gmat", MAXG * 9),
                    ("qacc_smooth", MAXNV), ("efc_force", MAXEFC)):
        d[name] = slice(o, o + n)
        o += n
    d["_size"] = o
    return d


class HostEmu:
    def __init__(self, blob: bytes, dtype=np.float64):
        self.L = lib()
        self.dt = 0 if dtype == np.float64 else 1
        self.np = np.float64 if self.dt == 0 else np.float32
        self.h = self.L.emu_create(blob, len(blob), self.dt)
        if not self.h:
            raise RuntimeError("emu_create failed")
        self.out_size = self.L.emu_out_size()
        assert self.out_size == out_slices()["_size"]

    def __del__(self):
        if getattr(self, "h", None):
            self.L.emu_destroy(self.h, self.dt)
            self.h = None

    def osc_config(self, gains6, null7, pt=5e-3, ot=68e-3):
        g, q = np.ascontiguousarray(gains6, np.float64), np.ascontiguousarray(null7, np.float64)
        self.L.emu_osc_config(self.h, self.dt, g.ctypes.data, q.ctypes.data, pt, ot)

    def run(self, mode, state, n_steps=1, control_steps=5, ctrl=None, target=None, grip=None, prop_size=None, prop_active=None,
            full_contacts=False, want_out=True, settle=(1e-3, 1e-2, 300), gains=None, solver=0, prop_type=None):
        """state: [n_env, stride] array of self.np (modified in place). Returns dict(out, status, converged, steps_done)."""
        assert state.dtype == self.np and state.flags.c_contiguous
        n = state.shape[0]
        keep = []

        def ptr(a, dt):
            if a is None:
                return None
            a = np.ascontiguousarray(a, dt)
            keep.append(a)
            return a.ctypes.data
        out = np.zeros((n, self.out_size), self.np) if want_out else None
        status, conv, done = np.zeros(n, np.int32), np.zeros(n, np.uint8), np.zeros(n, np.int32)
        a = EmuArgs(mode=mode, n_env=n, n_steps=n_steps, control_steps=control_steps, state_stride=state.shape[1], full_contacts=int(full_contacts),
                    settle_min_steps=settle[2], out_stride=self.out_size, settle_qvel_tol=settle[0], settle_qacc_tol=settle[1],
                    state=state.ctypes.data, ctrl=ptr(ctrl, self.np), target=ptr(target, self.np), grip=ptr(grip, np.uint8),
                    prop_size=ptr(prop_size, self.np), prop_active=ptr(prop_active, np.uint8), env_mask=None,
                    converged=conv.ctypes.data, status=status.ctypes.data, steps_done=done.ctypes.data, out=out.ctypes.data if want_out else None, gains=ptr(gains, self.np), solver=int(solver), prop_type=ptr(prop_type, np.uint8))
        self.L.emu_run(self.h, self.dt, C.byref(a))
        return dict(out=out, status=status, converged=conv, steps_done=done)
