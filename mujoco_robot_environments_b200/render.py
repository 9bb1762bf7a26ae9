"""BatchedRenderer: the reference's three ``mujoco.Renderer`` instances on the overhead camera (rgb, depth, segmentation;
tasks/rearrangement.py:182-186, :460-478, :273-275) for a batch of scenes, over the C-ABI of ``include/b2p_render.h``.

The reference rasterises with OpenGL (unavailable on this box); libb2p.so ray-casts the same scene on the GPU
(csrc/b2p_render.cu).  PyTorch only owns the output buffers.  There is no CPU path.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional

import torch

from . import ffi
from .physics import ASSETS, BatchedPhysics


def load_default_render_blob(tag: str = "rearrangement_n4") -> bytes:
    with open(os.path.join(ASSETS, tag + ".b2pr"), "rb") as f:
        return f.read()


class BatchedRenderer:
    def __init__(self, device=None, blob: Optional[bytes] = None):
        if not torch.cuda.is_available():
            raise RuntimeError("BatchedRenderer needs a CUDA device (B200); there is no CPU fallback")
        self.device = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        self._L = ffi.lib()
        self._h = C.c_void_p()
        blob = blob if blob is not None else load_default_render_blob()
        with torch.cuda.device(self.device):
            ffi.check(self._L.b2p_render_create(blob, len(blob), self.device.index or 0, C.byref(self._h)), "b2p_render_create")
        self.width, self.height = self.query("WIDTH"), self.query("HEIGHT")

    def query(self, what: str) -> int:
        return self._L.b2p_render_query(self._h, ffi.RQ[what])

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.b2p_render_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def alloc(self, n: int, seg: bool = True) -> Dict[str, torch.Tensor]:
        o = {"rgb": torch.empty(n, self.height, self.width, 3, dtype=torch.uint8, device=self.device),
             "depth": torch.empty(n, self.height, self.width, dtype=torch.float32, device=self.device)}
        if seg:
            o["seg"] = torch.empty(n, self.height, self.width, dtype=torch.int32, device=self.device)
        return o

    def render(self, physics: BatchedPhysics, prop_rgba: Optional[torch.Tensor] = None, out: Optional[Dict[str, torch.Tensor]] = None,
               lo: int = 0, hi: Optional[int] = None, seg: bool = True) -> Dict[str, torch.Tensor]:
        """Render scenes [lo, hi) of ``physics`` from the geom poses of its derived-output record (current after any launch;
        call ``physics.forward()`` after writing qpos by hand).  ``prop_rgba`` float32 [n_env, nprop, 4]."""
        hi = physics.n_env if hi is None else hi
        n = hi - lo
        out = out if out is not None else self.alloc(n, seg)
        es = physics.out.element_size()
        base = physics.out.data_ptr() + lo * physics._ostride * es
        xpos, xmat = base + physics._ofs["GEOM_XPOS"][0] * es, base + physics._ofs["GEOM_XMAT"][0] * es
        rgba = None
        if prop_rgba is not None:
            rgba = prop_rgba.to(device=self.device, dtype=torch.float32).contiguous()
            assert rgba.shape == (physics.n_env, physics.maxprop, 4)
        ps, pa = physics.prop_size, physics.prop_active
        self._keep = (rgba, out)
        ffi.check(self._L.b2p_render(self._h, n, xpos, xmat, physics._ostride, ffi.F64 if physics.dtype == torch.float64 else ffi.F32,
                                     None if rgba is None else rgba.data_ptr() + lo * physics.maxprop * 16,
                                     ps.data_ptr() + lo * physics.maxprop * 3 * ps.element_size(), pa.data_ptr() + lo * physics.maxprop,
                                     out["rgb"].data_ptr(), out["depth"].data_ptr(), out["seg"].data_ptr() if "seg" in out else None,
                                     C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)), "b2p_render")
        return out
