"""Host check of the ray-caster's per-tile plane lists (csrc/b2p_render_cull.h -- the very header the kernel includes): over random
poses of every hull of the render model, the clip loop over a tile's list returns the bits of the loop over all planes (hit / miss,
entry parameter, entry facet) on every pixel, and the lists are short enough to be worth having.  The GPU side of the same claim is
tests/test_gpu_render.py (bit-exact frames against the oracle, which knows nothing of tiles or lists)."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "hostemu", "b2p_render_cull_test.cpp")
HDR = os.path.join(ROOT, "mujoco_robot_environments_b200", "csrc", "b2p_render_cull.h")
EXE = os.path.join(ROOT, "tests", "_hostemu", "render_cull_test")
BLOB = os.path.join(ROOT, "mujoco_robot_environments_b200", "assets", "rearrangement_n4.b2pr")


def _build():
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    if not os.path.exists(EXE) or os.path.getmtime(EXE) < max(os.path.getmtime(SRC), os.path.getmtime(HDR)):
        # no FMA contraction, as in the kernel's translation unit (-fmad=false): the only fused operations are the explicit fmaf calls
        subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-o", EXE, SRC])
    return EXE


@pytest.mark.parametrize("seed", [1, 2])
def test_listed_clip_loop_is_bit_identical_to_the_full_loop(seed):
    out = subprocess.check_output([_build(), BLOB, str(seed), "10"], timeout=600)
    r = json.loads(out)
    assert r["rays"] > 2_000_000 and r["hits"] > 300_000, r
    assert r["mismatches"] == 0, r
    assert r["hit_kept"] < 0.4 * r["hit_planes"], r          # tiles that see the hull: about 30 % of its planes stay in the loop
    assert r["kept"] < 0.3 * r["planes"], r
