"""C-ABI checks that need no GPU: libb2p.so loads, exports every entry point include/b2p.h declares, and the
Python-side enum mirrors (ffi.Q / ffi.O) agree with the header.  No compute call is made here."""
import ctypes as C
import os
import re

import pytest

from mujoco_robot_environments_b200 import ffi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADERS = [os.path.join(ROOT, "include", "b2p.h"), os.path.join(ROOT, "include", "b2p_render.h")]


def _header_text():
    txt = ""
    for h in HEADERS:
        with open(h) as f:
            txt += f.read()
    return re.sub(r"/\*.*?\*/", "", txt, flags=re.S)


def _declared_functions():
    return sorted(set(re.findall(r"\b(b2p_[a-z0-9_]+)\s*\(", _header_text())))


def _enum_block(first_name):
    m = re.search(r"enum\s*\{([^}]*\b%s\b[^}]*)\}" % first_name, _header_text())
    names, val = [], 0
    for item in m.group(1).split(","):
        item = item.strip()
        if not item:
            continue
        if "=" in item:
            n, v = item.split("=")
            val = int(v.strip(), 0)
            item = n.strip()
        names.append((item, val))
        val += 1
    return dict(names)


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(ffi.LIB_PATH):
        ffi.build()
    return C.CDLL(ffi.LIB_PATH)


def test_header_declares_what_ffi_binds():
    assert _declared_functions() == sorted(ffi.SYMBOLS)


def test_library_exports_every_declared_symbol(lib):
    for name in _declared_functions():
        assert hasattr(lib, name), name


def test_enum_mirrors_match_header():
    q = _enum_block("B2P_Q_NQ")
    for k, v in ffi.Q.items():
        assert q["B2P_Q_" + k] == v, k
    o = _enum_block("B2P_O_SITE_XPOS")
    for k, v in ffi.O.items():
        assert o["B2P_O_" + k] == v, k
    assert o["B2P_O_NFIELD"] == len(ffi.O)
    r = _enum_block("B2P_RQ_WIDTH")
    for k, v in ffi.RQ.items():
        assert r["B2P_RQ_" + k] == v, k


def test_error_paths_without_a_gpu(lib):
    """Argument validation happens before any CUDA call: bad arguments return B2P_ERR_ARG (-1) and set last_error."""
    lib.b2p_last_error.restype = C.c_char_p
    h = C.c_void_p()
    assert lib.b2p_create(None, 0, 1, 0, 0, C.byref(h)) == -1
    assert b"b2p_create" in lib.b2p_last_error()
    assert lib.b2p_query(None, 0) == -1
    off, cnt = C.c_int(), C.c_int()
    assert lib.b2p_out_offset(9999, C.byref(off), C.byref(cnt)) == -1
    assert lib.b2p_out_offset(ffi.O["SITE_XPOS"], C.byref(off), C.byref(cnt)) == 0 and cnt.value == 6
    assert lib.b2p_step_physics(None, None, 1, None) == -1
    assert lib.b2p_destroy(None) == 0
    r = C.c_void_p()
    assert lib.b2p_render_create(None, 0, 0, C.byref(r)) == -1 and b"b2p_render_create" in lib.b2p_last_error()
    junk = bytes(4096)
    assert lib.b2p_render_create(junk, len(junk), 0, C.byref(r)) == -2          # not a B2PR blob: rejected before any CUDA call
    assert lib.b2p_render(None, 1, None, None, 0, 1, None, None, None, None, None, None, None) == -1
    assert lib.b2p_render_destroy(None) == 0


def test_bad_blob_is_rejected_before_touching_the_device(lib):
    h = C.c_void_p()
    junk = bytes(4096)
    rc = lib.b2p_create(junk, len(junk), 4, 0, 0, C.byref(h))
    assert rc == -2        # B2P_ERR_MODEL
    lib.b2p_last_error.restype = C.c_char_p
    assert b"magic" in lib.b2p_last_error()


def test_actuator_pattern_is_checked_at_create(lib):
    """the kernel folds osc.py:224-236 (tau @ pinv(actuator_moment)[arm, arm]) to the identity; b2p_create must refuse a
    model where that does not hold (geared arm motor, arm actuator on another dof, gripper actuator reaching an arm dof)."""
    from blobtools import surgery
    from mujoco_robot_environments_b200.physics import load_default_blob
    blob, _ = load_default_blob()
    lib.b2p_last_error.restype = C.c_char_p

    def geared(c, o, a): a["act_trncoef"][2, 0] = 2.0
    def moved(c, o, a): a["act_trndof"][3, 0] = 4
    def gripper_on_arm(c, o, a): a["act_trndof"][7, 1] = 6
    for edit in (geared, moved, gripper_on_arm):
        bad = surgery(blob, edit=edit)
        h = C.c_void_p()
        assert lib.b2p_create(bad, len(bad), 4, 0, 0, C.byref(h)) == -2, edit.__name__
        assert b"actuator" in lib.b2p_last_error()


def test_physics_backend_refuses_to_run_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from mujoco_robot_environments_b200.physics import BatchedPhysics
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        BatchedPhysics(4)
