"""Parity on BASELINE.json's own configurations: the GPU batch is driven exactly as bench.py drives it
(configs[1]: 4096 scenes, fp64, Halton torque noise; configs[2]: 16384 scenes, fp32, OSC in the loop), stopped, and a seeded
sample of 64 scenes -- always including the scenes with the most constraint rows and the most Newton iterations -- is replayed
on the GPU and on the CPU oracle: one step (qacc / qvel / qpos, contact sets) and 100 steps.  The same from states captured in
the middle of a scripted pick and place (gripper closing on the cube, cube in hand, gripper opening).

The MEASURED differences are written to gpurun_out/parity_*.json (copied to profiles/r02_parity_*.json and quoted in
DESIGN.md section 2); the asserts below are those measurements with head-room, against the north star's tolerances
(single step ~1e-6 relative in fp64, ~1e-4 in fp32; identical contact sets; bounded 100-step drift).
"""
import numpy as np
import pytest

import parity_lib as pl
import scenarios as sc

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

# north star: ~1e-6 relative (fp64), ~1e-4 relative (fp32) on a single step
TOL = {
    "f64": dict(qacc_rel=1e-6, qvel_rel=1e-9, qpos_rel=1e-12, drift100=1e-6, tau_rel=1e-9),
    "f32": dict(qacc_rel=1e-4, qvel_rel=1e-6, qpos_rel=1e-6, drift100=5e-3, tau_rel=1e-4),
}


def _check(name, r, min_identical=1.0, qacc_rel=None, drift=None):
    print(pl.summary_line(name, r))
    pl.save_report(name, r)
    t = TOL[r["dtype"]]
    s = r["single_step"]
    assert r["flags_single_step"] & 7 == 0 and r["flags_horizon"] & 7 == 0, "bad-state flag"
    assert s["qacc_rel"]["max"] < (qacc_rel or t["qacc_rel"]), s["qacc_rel"]
    assert s["qvel_rel"]["max"] < t["qvel_rel"], s["qvel_rel"]
    assert s["qpos_rel"]["max"] < t["qpos_rel"], s["qpos_rel"]
    assert r["contact_sets_identical"] >= min_identical * r["scenes"], (r["contact_sets_identical"], r["scenes"])
    last = r["qpos_drift"][max(r["qpos_drift"], key=int)]
    assert last["max"] < (drift or t["drift100"]), r["qpos_drift"]
    if "osc_tau_rel_max" in r:
        assert r["osc_tau_rel_max"] < t["tau_rel"], r["osc_tau_rel_max"]


def test_c2_batch_parity_fp64():
    """configs[1]: 4096 scenes, fp64, 500 steps of Halton torque noise, then 64 scenes against the oracle."""
    p, ctrl, size, active = pl.c2_batch(4096, launches=5, steps=100)
    assert int((p.status() & 7).max()) == 0
    idx = pl.batch_sample(p, 64, seed=0)
    st = pl.snapshot(p, idx)
    p.close()
    r = pl.compare_sample("f64", st, size[idx], active[idx], ctrl=ctrl[idx], horizon=100)
    _check("c2_f64", r)


def test_c3_batch_parity_fp32():
    """configs[2]: 16384 scenes, fp32, OSC in the loop, 400 control ticks (2000 steps), then 64 scenes against the oracle."""
    p, tg, grip, size, active = pl.c3_batch(16384, launches=20, ticks=20)
    assert int((p.status() & 7).max()) == 0
    idx = pl.batch_sample(p, 64, seed=0)
    st = pl.snapshot(p, idx)
    p.close()
    r = pl.compare_sample("f32", st, size[idx], active[idx], target=tg[idx], grip=grip[idx], horizon=100)
    _check("c3_f32", r, min_identical=0.9)


def test_c3_batch_parity_fp64():
    """the same OSC workload in the fp64 parity mode (2048 scenes): the kernel's OSC + step against the oracle at 1e-6"""
    p, tg, grip, size, active = pl.c3_batch(2048, launches=20, ticks=20)
    st = pl.snapshot(p, np.arange(2048)).astype(np.float32).astype(np.float64)      # fp32 states, replayed in fp64
    nefc = p.data.nefc.cpu().numpy(); niter = p.derived("NITER")[:, 0].cpu().numpy()
    idx = pl.pick_sample(nefc, niter, 64, seed=1)
    p.close()
    r = pl.compare_sample("f64", st[idx], size[idx], active[idx], target=tg[idx], grip=grip[idx], horizon=100)
    _check("c3_f64", r)


@pytest.mark.parametrize("dtype", ["f64", "f32"])
def test_pick_and_place_states_parity(dtype):
    """States captured mid-grasp, mid-lift (cube in hand) and mid-release of the scripted pick and place."""
    snaps, lifted = pl.pick_place_snapshots(dtype, n=32)
    assert lifted >= 24, lifted                          # the cube travels with the gripper in most scenes
    for name, (st, tg, grip, size, active) in snaps.items():
        r = pl.compare_sample(dtype, st, size, active, target=tg, grip=grip, horizon=100)
        assert r["nefc_max"] > 30
        _check("%s_%s" % (name, dtype), r, min_identical=0.9 if dtype == "f32" else 1.0)
