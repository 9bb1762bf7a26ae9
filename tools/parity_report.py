"""Measure (do not assert) the GPU-vs-oracle differences of tests/test_gpu_config_parity.py and print one summary line per case;
full per-scene rows go to gpurun_out/parity_<case>.json.   python tools/parity_report.py [case ...]"""
import os
import sys
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

import parity_lib as pl  # noqa: E402


def case_c2():
    p, ctrl, size, active = pl.c2_batch(4096, launches=5, steps=100)
    idx = pl.batch_sample(p, 64, seed=0); st = pl.snapshot(p, idx); p.close()
    return {"c2_f64": pl.compare_sample("f64", st, size[idx], active[idx], ctrl=ctrl[idx], horizon=100)}


def case_c3():
    p, tg, grip, size, active = pl.c3_batch(16384, launches=20, ticks=20)
    idx = pl.batch_sample(p, 64, seed=0); st = pl.snapshot(p, idx); p.close()
    return {"c3_f32": pl.compare_sample("f32", st, size[idx], active[idx], target=tg[idx], grip=grip[idx], horizon=100),
            "c3_f64_replay": pl.compare_sample("f64", st, size[idx], active[idx], target=tg[idx], grip=grip[idx], horizon=100)}


def case_pick(dtype):
    snaps, lifted = pl.pick_place_snapshots(dtype, n=32)
    out = {}
    for name, (st, tg, grip, size, active) in snaps.items():
        out["%s_%s" % (name, dtype)] = pl.compare_sample(dtype, st, size, active, target=tg, grip=grip, horizon=100)
        out["%s_%s" % (name, dtype)]["lifted"] = lifted
    return out


CASES = {"c2": case_c2, "c3": case_c3, "pick64": lambda: case_pick("f64"), "pick32": lambda: case_pick("f32")}

if __name__ == "__main__":
    for c in (sys.argv[1:] or list(CASES)):
        try:
            for name, r in CASES[c]().items():
                pl.save_report(name, r)
                print(pl.summary_line(name, r), flush=True)
        except Exception:
            traceback.print_exc()
