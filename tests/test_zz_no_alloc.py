"""b2p_step_physics_noalloc (include/b2p.h): the step kernel with its global-memory scratch in a file-scope static device array
instead of the handle's cudaMalloc'd pools.  No counterpart in the reference.

Checked through the C-ABI against the existing entry point, b2p_step_physics, on the same seeded scenes, on two calls in a row
(the static array keeps the first call's contents), within the tolerances of tests/test_gpu_parity.py's trajectory test.  The
two entry points run the same kernel on the same inputs, so in practice the states come out bitwise equal.

Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import scenarios as sc
from orc import Oracle

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

B2P_ERR_ARG = -1
MAX_ENV = 128                     # B2P_NOALLOC_MAX_ENV


def _setup(n, dtype, seed):
    from mujoco_robot_environments_b200.physics import BatchedPhysics
    st, size, active = sc.make_states(n, seed=seed, settled=False)       # props dropped 2 cm onto the table: contacts come and go
    p = BatchedPhysics(n, dtype=dtype)
    p.state.copy_(torch.tensor(st, dtype=dtype))
    p.set_env_params(torch.tensor(size), torch.tensor(active))
    o = Oracle(sc.default_blob())                                         # gravity compensation at the home pose, as in test_physics_steps_parity
    sc.oracle_load(o, st[0], size[0], active[0]); o.step1(); bias = o.qfrc_bias[:7].copy()
    p.set_control(torch.tensor(sc.halton_ctrl(n, 0, bias)))
    return p


def _stream(p):
    return C.c_void_p(torch.cuda.current_stream(p.device).cuda_stream)


@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-8), (torch.float32, 5e-3)])
def test_noalloc_step_matches_step_physics_twice_in_a_row(dtype, tol):
    n, steps = MAX_ENV, 60
    p = _setup(n, dtype, seed=21)
    L, h = p._L, p._h
    mine = p.state.clone()                                               # the caller's own state buffer for the new entry point
    assert mine.data_ptr() % 16 == 0
    for call in range(2):
        # existing entry point on the bound state
        rc = L.b2p_step_physics(h, p.ctrl.data_ptr(), steps, _stream(p)); assert rc == 0, L.b2p_last_error().decode()
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(p.device)[0]
        # new entry point on the caller's buffer
        rc = L.b2p_step_physics_noalloc(h, C.c_void_p(mine.data_ptr()), C.c_void_p(p.ctrl.data_ptr()), steps, _stream(p))
        assert rc == 0, L.b2p_last_error().decode()
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info(p.device)[0] == free0             # nothing was allocated or freed on the device
        want, got = p.state.cpu().numpy().astype(np.float64), mine.cpu().numpy().astype(np.float64)
        assert np.isfinite(got).all()
        assert np.abs(got[:, :sc.NQ] - want[:, :sc.NQ]).max() <= tol, (call, np.abs(got - want).max())
        assert np.abs(got - want).max() <= 100 * tol                     # velocities, warm start, time
        moved = np.abs(want[:, :sc.NQ] - sc.make_states(n, seed=21, settled=False)[0][:, :sc.NQ]).max()
        assert moved > 1e-3                                               # the scenes did advance


def test_noalloc_step_rejects_more_scenes_than_the_scratch_covers():
    p = _setup(MAX_ENV + 1, torch.float32, seed=22)
    L, h = p._L, p._h
    before = p.state.clone()
    rc = L.b2p_step_physics_noalloc(h, C.c_void_p(p.state.data_ptr()), None, 1, _stream(p))
    torch.cuda.synchronize()
    assert rc == B2P_ERR_ARG
    msg = L.b2p_last_error().decode()
    assert "B2P_NOALLOC_MAX_ENV" in msg and str(MAX_ENV) in msg, msg
    assert torch.equal(before, p.state)                                   # nothing ran
    # the existing entry point is untouched by the refusal
    assert L.b2p_step_physics(h, p.ctrl.data_ptr(), 1, _stream(p)) == 0
    torch.cuda.synchronize()
