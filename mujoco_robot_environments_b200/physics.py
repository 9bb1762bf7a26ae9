"""BatchedPhysics: the subset of ``dm_control.mjcf.Physics`` the reference's hot path uses, with a
leading env dimension, as torch tensors over the C-ABI of ``include/b2p.h``.

Reference surface mirrored: ``physics.step()`` / ``set_control`` (models/robot_arm.py:78-79),
``physics.forward()`` (environment/prop_initializer.py:190,221), ``physics.reset()`` (tasks/rearrangement.py:302),
``data.{time,qpos,qvel,site_xpos,contact}``, ``model.{nq,nv,nu,opt.timestep}`` and dm_control's scoped names
(``'panda nohand/robotiq_2f85/pinch'``, ``'table/table'``, ``'prop_0/prop_0'``).

PyTorch is plumbing here (device memory, streams); all arithmetic happens in libb2p.so.  There is no CPU path.
"""
from __future__ import annotations

import ctypes as C
import json
import os
from types import SimpleNamespace
from typing import Optional

import numpy as np
import torch

from . import ffi
from .compile import blob as B

ASSETS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "assets")
ARM_HOME = (0.0, -0.785, 0.0, -2.356, 0.0, 1.571, 0.785)   # config/robots/arm/franka_emika_panda.yaml:8


def load_default_blob(tag: str = "rearrangement_n4"):
    with open(os.path.join(ASSETS, tag + ".b2pm"), "rb") as f:
        blob = f.read()
    with open(os.path.join(ASSETS, tag + ".names.json")) as f:
        names = json.load(f)
    return blob, names


class BatchedPhysics:
    """``n_env`` independent rearrangement scenes stepped by the sm_100a kernels."""

    def __init__(self, n_env: int, dtype: torch.dtype = torch.float64, device: Optional[torch.device] = None,
                 blob: Optional[bytes] = None, names: Optional[dict] = None):
        if not torch.cuda.is_available():
            raise RuntimeError("BatchedPhysics needs a CUDA device (B200); there is no CPU fallback")
        if blob is None:
            blob, names = load_default_blob()
        self.device = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        self.dtype = dtype
        self.n_env = int(n_env)
        self.names = names or {}
        self._L = ffi.lib()
        self._h = C.c_void_p()
        dt = ffi.F64 if dtype == torch.float64 else ffi.F32
        with torch.cuda.device(self.device):
            ffi.check(self._L.b2p_create(blob, len(blob), self.n_env, dt, self.device.index or 0, C.byref(self._h)), "b2p_create")
        q = lambda k: self._L.b2p_query(self._h, ffi.Q[k])
        counts, opts, arrays = B.unpack(blob)
        self.model = SimpleNamespace(nq=q("NQ"), nv=q("NV"), nu=q("NU"), nprop=q("NPROP"), ngeom=q("NGEOM"),
                                     opt=SimpleNamespace(timestep=opts["timestep"], gravity=(opts["gravity_x"], opts["gravity_y"], opts["gravity_z"])),
                                     qpos0=torch.tensor(arrays["qpos0"], dtype=dtype), prop_qposadr=[int(x) for x in arrays["prop_qposadr"]],
                                     prop_dofadr=[int(x) for x in arrays["prop_dofadr"]], prop_geomid=[int(x) for x in arrays["prop_geomid"]],
                                     default_prop_size=arrays["geom_size"][arrays["prop_geomid"]].copy())
        self.maxprop, self.maxcon = q("MAXPROP"), q("MAXCON")
        self._stride, self._ostride = q("STATE_STRIDE"), q("OUT_STRIDE")
        self._rec = {k: q(k) for k in ("REC_QPOS", "REC_QVEL", "REC_QACC_WARMSTART", "REC_TIME")}
        kw = dict(device=self.device, dtype=dtype)
        # caller-owned buffers (the handle only borrows the pointers)
        self.state = torch.zeros(self.n_env, self._stride, **kw)
        self.out = torch.zeros(self.n_env, self._ostride, **kw)
        self.ctrl = torch.zeros(self.n_env, self.model.nu, **kw)
        self.prop_size = torch.tensor(self.model.default_prop_size, **kw).repeat(self.n_env, 1, 1).contiguous()
        if self.prop_size.shape[1] < self.maxprop:
            pad = torch.ones(self.n_env, self.maxprop - self.prop_size.shape[1], 3, **kw) * 0.0155
            self.prop_size = torch.cat([self.prop_size, pad], 1).contiguous()
        self.prop_active = torch.ones(self.n_env, self.maxprop, device=self.device, dtype=torch.uint8)
        self.prop_type = torch.full((self.n_env, self.maxprop), ffi.GEOM["box"], device=self.device, dtype=torch.uint8)   # collider class per prop
        self.converged = torch.zeros(self.n_env, device=self.device, dtype=torch.uint8)
        self.steps_done = torch.zeros(self.n_env, device=self.device, dtype=torch.int32)
        self._status = torch.zeros(self.n_env, device=self.device, dtype=torch.int32)
        ffi.check(self._L.b2p_bind_state(self._h, self.state.data_ptr()), "b2p_bind_state")
        ffi.check(self._L.b2p_bind_out(self._h, self.out.data_ptr()), "b2p_bind_out")
        ffi.check(self._L.b2p_set_env_params(self._h, self.prop_size.data_ptr(), self.prop_active.data_ptr()), "b2p_set_env_params")
        self._classes_bound = False        # the class-capable kernels are selected only once a scene holds a sphere / cylinder prop
        self._ofs = {}
        for k, fid in ffi.O.items():
            o, c = C.c_int(), C.c_int()
            ffi.check(self._L.b2p_out_offset(fid, C.byref(o), C.byref(c)))
            self._ofs[k] = (o.value, c.value)
        self.data = _Data(self)
        self.out_mask = ffi.OUT["ALL"]
        self.reset()

    # ------------------------------------------------------------------ lifetime
    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.b2p_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # ------------------------------------------------------------------ state views (no copies)
    @property
    def qpos(self): return self.state[:, self._rec["REC_QPOS"]:self._rec["REC_QPOS"] + self.model.nq]
    @property
    def qvel(self): return self.state[:, self._rec["REC_QVEL"]:self._rec["REC_QVEL"] + self.model.nv]
    @property
    def qacc_warmstart(self): return self.state[:, self._rec["REC_QACC_WARMSTART"]:self._rec["REC_QACC_WARMSTART"] + self.model.nv]
    @property
    def time(self): return self.state[:, self._rec["REC_TIME"]]

    def derived(self, field: str) -> torch.Tensor:
        o, c = self._ofs[field]
        return self.out[:, o:o + c]

    # ------------------------------------------------------------------ dm_control-like surface
    def reset(self, mask: Optional[torch.Tensor] = None):
        """physics.reset(): qpos = qpos0, everything else zero (mj_resetData)."""
        rec = torch.zeros(self._stride, device=self.device, dtype=self.dtype)
        rec[self._rec["REC_QPOS"]:self._rec["REC_QPOS"] + self.model.nq] = self.model.qpos0.to(self.device)
        if mask is None:
            self.state[:] = rec
            self.ctrl.zero_()
        else:
            self.state[mask] = rec
            self.ctrl[mask] = 0

    def set_control(self, ctrl: torch.Tensor):
        self.ctrl.copy_(ctrl.to(self.ctrl))

    def step(self, n_steps: int = 1):
        """n x (physics.set_control(ctrl); physics.step()) with ctrl held (robot_arm.py:77-79)."""
        ffi.check(self._L.b2p_step_physics(self._h, self.ctrl.data_ptr(), int(n_steps), self._stream()), "b2p_step_physics")

    def forward(self, full_contacts: bool = True):
        """physics.forward(): derived quantities for the current state; contacts listed inside geom margin."""
        ffi.check(self._L.b2p_forward(self._h, int(full_contacts), self._stream()), "b2p_forward")

    def run_controller(self, target13: torch.Tensor, grip: torch.Tensor, n_ticks: int, control_steps: int = 5) -> torch.Tensor:
        """RobotArm.run_controller(duration) for every env: n_ticks x (OSC + gripper command + control_steps steps)."""
        t = target13.to(device=self.device, dtype=self.dtype).contiguous()
        g = grip.to(device=self.device, dtype=torch.uint8).contiguous()
        assert t.shape == (self.n_env, 13) and g.shape == (self.n_env,)
        ffi.check(self._L.b2p_step_osc(self._h, t.data_ptr(), g.data_ptr(), int(n_ticks), int(control_steps), self.converged.data_ptr(), self._stream()), "b2p_step_osc")
        self._keep = (t, g)
        return self.converged

    def settle(self, max_steps: int = 2000, min_steps: int = 300, qvel_tol: float = 1e-3, qacc_tol: float = 1e-2, mask: Optional[torch.Tensor] = None):
        """PropPlacer settle loop with the arm held static (prop_initializer.py:245-257)."""
        m = None if mask is None else mask.to(device=self.device, dtype=torch.uint8).contiguous()
        self._keep = m
        ffi.check(self._L.b2p_settle(self._h, None if m is None else m.data_ptr(), int(max_steps), int(min_steps), float(qvel_tol), float(qacc_tol),
                                     self.steps_done.data_ptr(), self._stream()), "b2p_settle")
        return self.steps_done

    def place_search(self, prop_id: torch.Tensor, lo: torch.Tensor, hi: torch.Tensor, table_geom: int, dist: float, max_attempts: int,
                     seed: int, mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        """SortColours.prop_place for every scene in one launch (``b2p_place_search``) -> [n_env, 4] = accepted xyz, attempts used
        (negative: nothing accepted).  State and derived outputs are untouched; rows outside ``mask`` are zero."""
        pid = prop_id.to(device=self.device, dtype=torch.int32).contiguous()
        box = torch.cat([lo, hi], 1).to(device=self.device, dtype=self.dtype).contiguous()
        m = None if mask is None else mask.to(device=self.device, dtype=torch.uint8).contiguous()
        assert pid.shape == (self.n_env,) and box.shape == (self.n_env, 6)
        out = torch.zeros(self.n_env, 4, device=self.device, dtype=self.dtype)
        ffi.check(self._L.b2p_place_search(self._h, pid.data_ptr(), box.data_ptr(), None if m is None else m.data_ptr(), int(max_attempts), float(dist),
                                           int(table_geom), int(seed) & 0xFFFFFFFF, out.data_ptr(), self._stream()), "b2p_place_search")
        self._keep = (pid, box, m)
        return out

    def place_props(self, prop_id: torch.Tensor, lo: torch.Tensor, hi: torch.Tensor, table_geom: int, dist: float, max_attempts: int,
                    seed: int, mask: Optional[torch.Tensor] = None):
        """PropPlacer for one prop per scene in one launch (``b2p_place_props``): proposals = position ~ U(lo, hi), yaw = pi * U about z
        -> (out [n_env, 4] = accepted xyz, attempts used (negative: none accepted), quat [n_env, 4]).  The state is not written."""
        pid = prop_id.to(device=self.device, dtype=torch.int32).contiguous()
        box = torch.cat([lo, hi], 1).to(device=self.device, dtype=self.dtype).contiguous()
        m = None if mask is None else mask.to(device=self.device, dtype=torch.uint8).contiguous()
        assert pid.shape == (self.n_env,) and box.shape == (self.n_env, 6)
        out = torch.zeros(self.n_env, 4, device=self.device, dtype=self.dtype)
        quat = torch.zeros(self.n_env, 4, device=self.device, dtype=self.dtype)
        ffi.check(self._L.b2p_place_props(self._h, pid.data_ptr(), box.data_ptr(), None if m is None else m.data_ptr(), int(max_attempts), float(dist),
                                          int(table_geom), int(seed) & 0xFFFFFFFF, out.data_ptr(), quat.data_ptr(), self._stream()), "b2p_place_props")
        self._keep = (pid, box, m)
        return out, quat

    def osc_config(self, gains6, nullspace7, pos_thresh=5e-3, ori_thresh=68e-3):
        g = (C.c_double * 6)(*[float(x) for x in gains6])
        q = (C.c_double * 7)(*[float(x) for x in nullspace7])
        ffi.check(self._L.b2p_osc_config(self._h, g, q, float(pos_thresh), float(ori_thresh)), "b2p_osc_config")

    def set_env_gains(self, gains: Optional[torch.Tensor]):
        """Per-scene OSC gains [n_env, 6] (position kp, kd, orientation kp, kd, nullspace kp, kd), or None for the handle's."""
        if gains is None:
            self._env_gains = None
            ffi.check(self._L.b2p_set_env_gains(self._h, None))
            return
        g = gains.to(device=self.device, dtype=self.dtype).contiguous()
        assert g.shape == (self.n_env, 6)
        self._env_gains = g                      # borrowed by the handle: keep it alive
        ffi.check(self._L.b2p_set_env_gains(self._h, g.data_ptr()))

    def set_env_mask(self, mask: Optional[torch.Tensor]):
        """Sticky scene mask for ``step`` / ``run_controller`` (bool / uint8 [n_env]; None = all): masked-out scenes are not
        stepped and keep their state, time and ``converged`` bit -- their episode is over."""
        if mask is None:
            self._env_mask = None
            ffi.check(self._L.b2p_set_env_mask(self._h, None))
            return
        m = mask.to(device=self.device, dtype=torch.uint8).contiguous()
        assert m.shape == (self.n_env,)
        self._env_mask = m                       # borrowed by the handle: keep it alive
        ffi.check(self._L.b2p_set_env_mask(self._h, m.data_ptr()))

    def set_out_fields(self, *groups: str):
        """Which groups of the derived-output record the stepping calls refresh (``b2p_set_out_mask``): any of "SITE", "CONTACT",
        "MARM", "GEOM", "BIAS", "DYN", "EFC", or "ALL" (the default).  ``forward()`` always refreshes everything.  Groups left out
        keep the values of the last launch that wrote them -- a caller that reads ``data.contact`` after ``step()`` must keep
        "CONTACT" in (or call ``forward()`` first, as the reference does before it scans contacts)."""
        mask = 0
        for g in groups:
            mask |= ffi.OUT[g.upper()]
        ffi.check(self._L.b2p_set_out_mask(self._h, mask), "b2p_set_out_mask")
        self.out_mask = mask

    def set_solver(self, name: str):
        """mjOption.solver for the stepping calls: "newton" (MuJoCo's and this backend's default) or "pgs" (``b2p_set_solver``)."""
        ffi.check(self._L.b2p_set_solver(self._h, ffi.SOLVER[name.lower()]), "b2p_set_solver")

    def set_full_contacts(self, full: bool):
        ffi.check(self._L.b2p_set_full_contacts(self._h, int(full)))

    def set_env_params(self, prop_size: Optional[torch.Tensor] = None, prop_active: Optional[torch.Tensor] = None,
                       prop_type: Optional[torch.Tensor] = None):
        """Per-scene prop parameters: half sizes [n, P, 3] (sphere: [r, -, -]; cylinder: [r, half height, -]), active flags [n, P] and
        collider class [n, P] (``ffi.GEOM``: box 6 / sphere 2 / cylinder 5; environment/props.py draws a shape per prop)."""
        if prop_type is not None:
            self.prop_type.copy_(prop_type.to(self.prop_type))
            want = bool((self.prop_type != ffi.GEOM["box"]).any())
            if want != self._classes_bound:       # box-only batches keep the throughput kernel (b2p_set_env_prop_types(NULL))
                ffi.check(self._L.b2p_set_env_prop_types(self._h, self.prop_type.data_ptr() if want else None), "b2p_set_env_prop_types")
                self._classes_bound = want
        if prop_size is not None:
            self.prop_size.copy_(prop_size.to(self.prop_size))
        if prop_active is not None:
            self.prop_active.copy_(prop_active.to(self.prop_active))

    # bit k of status(): the first three are mj_checkPos / Vel / Acc, the rest capacity / iteration-cap notes (csrc/b2p_math.h)
    STATUS_FLAGS = ("badqpos", "badqvel", "badqacc", "contactfull", "efcfull", "candfull", "solver_maxiter")

    def status(self, clear=False) -> torch.Tensor:
        """Per-env OR of bad-state flags (dm_control check_invalid_state equivalent).  ``clear``: True clears every scene's flags after
        the read, a bool tensor [n_env] clears those scenes' only (a masked reset must not wipe what the other scenes raised)."""
        if isinstance(clear, torch.Tensor):
            ffi.check(self._L.b2p_status(self._h, self._status.data_ptr(), 0, self._stream()))
            kept = torch.where(clear.to(self.device, torch.bool), torch.zeros_like(self._status), self._status)
            ffi.check(self._L.b2p_status_set(self._h, kept.data_ptr(), self._stream()))
            self._keep_status = kept
            return self._status
        ffi.check(self._L.b2p_status(self._h, self._status.data_ptr(), int(clear), self._stream()))
        return self._status

    @property
    def launches(self) -> int:
        return self._L.b2p_query(self._h, ffi.Q["LAUNCHES"])

    @property
    def order_launches(self) -> int:
        """Launches of the scene-queue ordering kernel (csrc/b2p_api.cu: b2p_order_kernel)."""
        return self._L.b2p_query(self._h, ffi.Q["ORDER_LAUNCHES"])

    def launch_info(self) -> dict:
        return {k.lower(): self._L.b2p_query(self._h, ffi.Q[k]) for k in ("SMEM_BYTES", "BLOCKS_PER_SM", "THREADS", "GRID")}

    # names (dm_control scoping)
    def geom_id(self, name: str) -> int:
        return self.names["geom"].index(name)


class _Data:
    """``physics.data``-like accessor: tensors with a leading env dim."""

    def __init__(self, p: BatchedPhysics):
        self._p = p

    @property
    def qpos(self): return self._p.qpos
    @property
    def qvel(self): return self._p.qvel
    @property
    def time(self): return self._p.time
    @property
    def ctrl(self): return self._p.ctrl
    @property
    def qacc(self): return self._p.derived("QACC")[:, :self._p.model.nv]
    @property
    def qfrc_bias(self): return self._p.derived("QFRC_BIAS")[:, :self._p.model.nv]
    @property
    def site_xpos(self): return self._p.derived("SITE_XPOS").view(self._p.n_env, 2, 3)
    @property
    def site_xmat(self): return self._p.derived("SITE_XMAT").view(self._p.n_env, 2, 9)
    @property
    def geom_xpos(self): return self._p.derived("GEOM_XPOS").reshape(self._p.n_env, -1, 3)[:, :self._p.model.ngeom]
    @property
    def geom_xmat(self): return self._p.derived("GEOM_XMAT").reshape(self._p.n_env, -1, 9)[:, :self._p.model.ngeom]
    @property
    def ncon(self): return self._p.derived("NCON")[:, 0].to(torch.int64)
    @property
    def nefc(self): return self._p.derived("NEFC")[:, 0].to(torch.int64)

    @property
    def contact(self):
        """dict(dist [B,C], pos [B,C,3], normal [B,C,3], geom1, geom2, exclude) padded to MAXCON; valid where index < ncon."""
        c = self._p.derived("CONTACT").reshape(self._p.n_env, self._p.maxcon, 10)
        return dict(dist=c[..., 0], pos=c[..., 1:4], normal=c[..., 4:7], geom1=c[..., 7].to(torch.int64), geom2=c[..., 8].to(torch.int64),
                    exclude=c[..., 9].to(torch.int64))
