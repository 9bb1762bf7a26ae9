"""ctypes wrapper around the CPU oracle (oracle/libb2p_oracle.so). Test infrastructure only."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_FIELDS = ["QPOS", "QVEL", "CTRL", "QACC_WARMSTART", "TIME", "QACC", "QACC_SMOOTH", "QFRC_BIAS", "QFRC_PASSIVE",
           "QFRC_ACTUATOR", "QFRC_SMOOTH", "QFRC_CONSTRAINT", "XPOS", "XQUAT", "XMAT", "XIPOS", "GEOM_XPOS", "GEOM_XMAT",
           "SITE_XPOS", "SITE_XMAT", "QM", "SUBTREE_COM", "CINERT", "CDOF", "CVEL", "CDOF_DOT", "EFC_J", "EFC_POS", "EFC_D",
           "EFC_R", "EFC_AREF", "EFC_FORCE", "EFC_VEL", "EFC_MARGIN", "PROP_SIZE", "ACT_FORCE", "OSC_TAU", "SOLVER_STATS"]
F = {n: i for i, n in enumerate(_FIELDS)}
_lib = None


def lib():
    global _lib
    if _lib is None:
        # ORC_FAST=1 selects the timing build (-O3 -march=native; bench.py's CPU arm); parity tests use the strict build
        so = os.path.join(ROOT, "oracle", "libb2p_oracle_fast.so" if os.environ.get("ORC_FAST") == "1" else "libb2p_oracle.so")
        src = os.path.join(ROOT, "oracle", "b2p_oracle.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL)
        L = C.CDLL(so)
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.c_char_p, C.c_size_t]
        L.orc_ptr.restype = C.POINTER(C.c_double)
        L.orc_ptr.argtypes = [C.c_void_p, C.c_int]
        for fn in ("orc_destroy", "orc_reset", "orc_step1", "orc_step2", "orc_step", "orc_forward", "orc_step_frozen_arm"):
            getattr(L, fn).argtypes = [C.c_void_p]
            getattr(L, fn).restype = None
        L.orc_dim.argtypes = [C.c_void_p, C.c_int]
        L.orc_set_prop.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double), C.c_int]
        L.orc_set_prop_type.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.orc_set_option.argtypes = [C.c_void_p, C.c_int, C.c_double]
        L.orc_contacts.argtypes = [C.c_void_p] + [C.c_void_p] * 5
        L.orc_osc_config.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double]
        L.orc_osc_torque.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_osc_converged.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_run_controller.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
        _lib = L
    return _lib


class Oracle:
    DIMS = dict(nq=0, nv=1, nu=2, nbody=3, ngeom=4, nsite=5, ncon=6, nefc=7, nprop=8, solver_iter=9)

    def __init__(self, blob: bytes):
        self.L = lib()
        self.h = self.L.orc_create(blob, len(blob))
        if not self.h:
            raise RuntimeError("orc_create failed")
        for k, i in self.DIMS.items():
            if k not in ("ncon", "nefc", "solver_iter"):
                setattr(self, k, self.L.orc_dim(self.h, i))
        self._shape = dict(QPOS=(self.nq,), QVEL=(self.nv,), CTRL=(self.nu,), QACC_WARMSTART=(self.nv,), TIME=(1,),
                           QACC=(self.nv,), QACC_SMOOTH=(self.nv,), QFRC_BIAS=(self.nv,), QFRC_PASSIVE=(self.nv,),
                           QFRC_ACTUATOR=(self.nv,), QFRC_SMOOTH=(self.nv,), QFRC_CONSTRAINT=(self.nv,),
                           XPOS=(self.nbody, 3), XQUAT=(self.nbody, 4), XMAT=(self.nbody, 9), XIPOS=(self.nbody, 3),
                           GEOM_XPOS=(self.ngeom, 3), GEOM_XMAT=(self.ngeom, 9), SITE_XPOS=(self.nsite, 3),
                           SITE_XMAT=(self.nsite, 9), QM=(self.nv, self.nv), SUBTREE_COM=(self.nbody, 3),
                           CINERT=(self.nbody, 10), CDOF=(self.nv, 6), CVEL=(self.nbody, 6), CDOF_DOT=(self.nv, 6),
                           PROP_SIZE=(self.nprop, 3), ACT_FORCE=(self.nu,), OSC_TAU=(7,), SOLVER_STATS=(4,))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_destroy(self.h)
            self.h = None

    def dim(self, k):
        return self.L.orc_dim(self.h, self.DIMS[k])

    def view(self, name):
        """numpy view (no copy) into oracle memory."""
        p = self.L.orc_ptr(self.h, F[name])
        if name.startswith("EFC_"):
            n = self.dim("nefc")
            shape = (n, self.nv) if name == "EFC_J" else (n,)
        else:
            shape = self._shape[name]
        n = int(np.prod(shape))
        if n == 0:
            return np.zeros(shape)
        return np.ctypeslib.as_array(p, shape=(n,)).reshape(shape)

    def __getattr__(self, k):
        if k.upper() in F:
            return self.view(k.upper())
        raise AttributeError(k)

    def reset(self): self.L.orc_reset(self.h)
    def step1(self): self.L.orc_step1(self.h)
    def step2(self): self.L.orc_step2(self.h)
    def step(self): self.L.orc_step(self.h)
    def forward(self): self.L.orc_forward(self.h)
    def step_frozen_arm(self): self.L.orc_step_frozen_arm(self.h)

    def set_prop(self, p, half, active=True):
        a = (C.c_double * 3)(*[float(x) for x in half])
        self.L.orc_set_prop(self.h, p, a, int(active))

    def set_prop_type(self, p, geom_type):
        """re-class prop p's collider: blob.GEOM_BOX (as compiled) / GEOM_SPHERE / GEOM_CYLINDER (include/b2p_prop_types.h)"""
        self.L.orc_set_prop_type(self.h, int(p), int(geom_type))

    def set_option(self, what, v):
        self.L.orc_set_option(self.h, dict(full_contacts=0, integrator=1, tolerance=2, iterations=3, solver=4)[what], float(v))

    def contacts(self):
        n = self.dim("ncon")
        dist, pos, frame = np.zeros(n), np.zeros((n, 3)), np.zeros((n, 9))
        geom, excl = np.zeros((n, 2), np.int32), np.zeros(n, np.int32)
        self.L.orc_contacts(self.h, dist.ctypes.data, pos.ctypes.data, frame.ctypes.data, geom.ctypes.data, excl.ctypes.data)
        return dict(dist=dist, pos=pos, frame=frame, geom=geom, exclude=excl)

    def osc_config(self, gains6, null7, pos_thresh=5e-3, ori_thresh=68e-3):
        g, q = np.ascontiguousarray(gains6, float), np.ascontiguousarray(null7, float)
        self.L.orc_osc_config(self.h, g.ctypes.data, q.ctypes.data, pos_thresh, ori_thresh)

    def osc_torque(self, target13):
        t, tau = np.ascontiguousarray(target13, float), np.zeros(7)
        self.L.orc_osc_torque(self.h, t.ctypes.data, tau.ctypes.data)
        return tau

    def osc_converged(self, target13):
        t = np.ascontiguousarray(target13, float)
        return bool(self.L.orc_osc_converged(self.h, t.ctypes.data))

    def run_controller(self, target13, grip_max, n_ticks, control_steps=5):
        t = np.ascontiguousarray(target13, float)
        return bool(self.L.orc_run_controller(self.h, t.ctypes.data, int(grip_max), int(n_ticks), int(control_steps)))
