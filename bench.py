"""bench.py -- env-steps/sec (physics+OSC) of the batched rearrangement hot path on N B200s.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3|c2] [--envs E] [--ticks T] [--impl reference]

One "step" = one launch of the hot path over the whole env batch:
  c3 (default; BASELINE.json configs[2], the config the metric "env-steps/sec (physics+OSC)" is quoted on):
      16384 envs per GPU, fp32, colour_splitter props (N ~ U{2,3,4}, padded to 4), OSC in the loop,
      T control ticks x 5 physics steps per launch (RobotArm.run_controller semantics);
  c2 (configs[1]): 4096 envs, fp64, physics only under Halton control noise, 5*T steps per launch.
Rank 0 prints ONE JSON line (see the task contract).  `--impl reference` times the CPU restatement of the
reference path (oracle; MuJoCo itself is not installable here) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import multiprocessing as mp
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {
    "c3": dict(envs=16384, dtype="f32", osc=True, desc="colour_splitter props + OSC in the loop, 16384 envs/GPU, fp32 (BASELINE configs[2])"),
    "c2": dict(envs=4096, dtype="f64", osc=False, desc="default rearrangement scene physics-only, 4096 envs, fp64 (BASELINE configs[1])"),
    # side workload: overhead-camera observations of configs[3]; metric frames/s, HBM-write bound
    "render": dict(envs=2048, dtype="f32", osc=False, desc="overhead camera rgb+depth+segmentation 480x640 of 2048 scenes/GPU, arm in view in 1 of 4 scenes"),
}
CONTROL_STEPS = 5
# per-scene status bits (csrc/b2p_math.h): the first three are mj_checkPos/Vel/Acc; the rest are capacity / iteration-cap notes
FLAG_NAMES = ["badqpos", "badqvel", "badqacc", "contactfull", "efcfull", "candfull", "solver_maxiter"]


def measured_traffic(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the step kernel from the committed ncu --set full capture."""
    try:
        return float(json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[workload]["dram_bytes_per_launch"])
    except Exception:
        return None


# ------------------------------------------------------------------------------------------ CPU baseline (oracle port)
def _cpu_worker(args):
    seed, seconds, osc = args
    import numpy as np
    import scenarios as sc
    from orc import Oracle
    o = Oracle(sc.default_blob())
    st, size, active = sc.make_states(1, seed=1000 + seed)
    sc.oracle_load(o, st[0], size[0], active[0]); o.set_option("full_contacts", 0)
    tg, grip = sc.random_targets(64, seed=seed)
    o.step1(); bias = o.qfrc_bias[:7].copy()
    steps, k, t0 = 0, 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        if osc:
            o.run_controller(tg[k % 64], int(grip[k % 64]), 20, CONTROL_STEPS); steps += 20 * CONTROL_STEPS
        else:
            o.ctrl[:] = sc.halton_ctrl(1, k, bias)[0]
            for _ in range(100):
                o.step()
            steps += 100
        k += 1
    return steps, time.perf_counter() - t0


CPU_BUILD = "gcc -O3 -march=native -DNDEBUG -DORC_FAST (oracle/Makefile: libb2p_oracle_fast.so; hill-climb hull support, OBB mid phase, sparse rows, no per-step malloc)"


def cpu_baseline(seconds: float, osc: bool, cores: int | None = None):
    """The CPU arm: the oracle's TIMING build (same source as the strict parity build, same results), one scene per process on every host core."""
    cores = cores or os.cpu_count() or 1
    os.environ["ORC_FAST"] = "1"                      # inherited by the spawned workers: tests/orc.py loads libb2p_oracle_fast.so
    # -march=native: always rebuilt on the machine that times it (the .so in the snapshot was built on another CPU)
    subprocess.check_call(["make", "-B", "-C", os.path.join(ROOT, "oracle"), "libb2p_oracle_fast.so"], stdout=subprocess.DEVNULL)
    with mp.get_context("spawn").Pool(cores) as pool:
        res = pool.map(_cpu_worker, [(i, seconds, osc) for i in range(cores)])
    return sum(s / t for s, t in res), cores


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.rows = index, False, []

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        mhz = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None, "sm_max_mhz": int(self.rows[0][1]) if self.rows[0][1].isdigit() else None, "reasons": reasons}


# ------------------------------------------------------------------------------------------ arms
def workload_config(args, wl):
    """the keys that NAME the workload: identical in both arms (the driver compares them)"""
    return {"workload": args.workload + ": " + wl["desc"], "envs_per_gpu": args.envs or wl["envs"], "control_ticks_per_launch": args.ticks,
            "physics_steps_per_launch": args.ticks * CONTROL_STEPS}


def run_reference(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    per_step_s = max(2.0, min(20.0, 60.0 / max(1, args.steps + args.warmup)))
    vals = []
    for i in range(args.warmup + args.steps):
        v, cores = cpu_baseline(per_step_s, wl["osc"])
        if i >= args.warmup:
            vals.append(v)
    value = sum(vals) / len(vals)
    sample = "%d processes x %.1f s each of the same scenes/targets, one env per process (oracle C port of mj_step + OSC, %s; MuJoCo not installable)" % (cores, per_step_s, CPU_BUILD)
    line = {"impl": "reference", "metric": "env-steps/sec (physics+OSC)" if wl["osc"] else "env-steps/sec (physics only)", "value": value, "unit": "env-steps/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": per_step_s * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(args, wl),
            "cpu_baseline": {"value": value, "unit": "env-steps/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def _render_cpu_worker(args):
    seed, seconds = args
    import numpy as np
    import rnd
    import scenarios as sc
    from orc import Oracle
    st, size, active = sc.make_states(1, seed=seed)
    o = Oracle(sc.default_blob()); sc.oracle_load(o, st[0], size[0], active[0]); o.forward()
    blob = rnd.render_blob()
    frames, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        rnd.render(blob, o.geom_xpos, o.geom_xmat, None, size[0], active[0]); frames += 1
    return frames, time.perf_counter() - t0


def run_render(args, wl):
    """frames/s of b2p_render (rgb u8 + depth f32 + seg i32 per frame = 11 B/pixel written).  One "step" = one frame of every scene."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import scenarios as sc
    from mujoco_robot_environments_b200.physics import BatchedPhysics
    from mujoco_robot_environments_b200.render import BatchedRenderer
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local); dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = args.envs or wl["envs"]
    st, size, active = sc.make_states(n, seed=1234 + rank)
    st[::4, :7] = np.asarray([0.0, 0.35, 0.0, -1.9, 0.0, 2.25, 0.785]) + 0.2 * (np.random.default_rng(7).random((len(st[::4]), 7)) - 0.5)
    p = BatchedPhysics(n, dtype=torch.float32, device=dev)
    p.state.copy_(torch.tensor(st, dtype=torch.float32)); p.set_env_params(torch.tensor(size), torch.tensor(active)); p.forward(full_contacts=False)
    r = BatchedRenderer(dev)
    rgba = torch.tensor(np.random.default_rng(1).uniform(0, 1, (n, 4, 4)).astype(np.float32), device=dev)
    out = r.alloc(n)
    H, W = r.height, r.width
    bytes_per_frame = H * W * 11
    h_rgb = torch.empty(min(n, 64), H, W, 3, dtype=torch.uint8).pin_memory()
    for _ in range(args.warmup):
        r.render(p, rgba, out=out)
    torch.cuda.synchronize()
    sampler = ClockSampler(local); sampler.start()
    l0 = r.query("LAUNCHES")
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    for i in range(args.steps):             # outputs (n x 3.4 MB >> 126 MB L2) are larger than L2: no flush needed
        ev[i][0].record(); r.render(p, rgba, out=out); ev[i][1].record()
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in ev)
    launches = r.query("LAUNCHES") - l0
    # end to end: poses come from a physics launch on host-provided state; a batch of frames is read back to pinned host memory
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h_state = torch.tensor(st, dtype=torch.float32).pin_memory()
    e0.record()
    for i in range(max(2, args.steps)):
        p.state.copy_(h_state, non_blocking=True); p.forward(full_contacts=False); r.render(p, rgba, out=out); h_rgb.copy_(out["rgb"][:h_rgb.shape[0]], non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    e2e_ms = e0.elapsed_time(e1) / max(2, args.steps)
    sampler.stop_flag = True; sampler.join(timeout=2)
    t = torch.tensor([ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        value = n * world * args.steps / (t[0].item() * 1e-3)
        try:
            peak, which = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            peak, which = 6650.0, "fallback (B200_PROFILING.md)"
        achieved = bytes_per_frame * n / (t[0].item() * 1e-3 / args.steps) / 1e9
        copied, cast = r.query("TILES_COPIED"), r.query("TILES_CAST")
        line = {"metric": "overhead-camera frames/sec (rgb+depth+seg)", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": t[0].item() / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": "render: " + wl["desc"], "envs_per_gpu": n, "l2": "outputs (%.1f GB per step) exceed L2" % (bytes_per_frame * n / 1e9),
                           "tiles_copied_frac": copied / max(1, copied + cast)},
                "clocks": sampler.summary(), "gpu_launches": int(launches) * world,
                "e2e": {"value": n * world / (t[1].item() * 1e-3), "unit": "frames/s", "h2d_bytes_per_step": int(h_state.numel() * 4), "d2h_bytes_per_step": int(h_rgb.numel())},
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": measured_traffic("render"),
                             "peak_source": which, "algorithmic_bytes_per_frame": bytes_per_frame}}
        if world == 1 and not args.no_cpu:
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL)
            cores = os.cpu_count() or 1
            with mp.get_context("spawn").Pool(cores) as pool:
                res = pool.map(_render_cpu_worker, [(i, args.cpu_seconds) for i in range(cores)])
            line["cpu_baseline"] = {"value": sum(f / s for f, s in res), "unit": "frames/s", "cores": cores, "kind": "port",
                                    "sample": "%d processes x %.0f s of full 480x640 frames (plain-C brute-force restatement; GL renderer unavailable on this box)" % (cores, args.cpu_seconds)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def measure(args, wl, workload, n_steps):
    """Times `n_steps` launches of one workload on this rank's GPU (+ the end-to-end leg); rank 0 gets the JSON line, others None."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import scenarios as sc
    from mujoco_robot_environments_b200.physics import BatchedPhysics

    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    n = (args.envs if workload == args.workload else 0) or wl["envs"]
    args = argparse.Namespace(**{**vars(args), "steps": n_steps, "workload": workload})
    dtype = torch.float32 if wl["dtype"] == "f32" else torch.float64
    w = 4 if wl["dtype"] == "f32" else 8
    ticks = args.ticks
    steps_per_launch = ticks * CONTROL_STEPS
    # synthetic scenes (seed 1234 + rank), settled on device before timing
    st, size, active = sc.make_states(n, seed=1234 + rank, settled=True)
    p = BatchedPhysics(n, dtype=dtype, device=dev)
    p.state.copy_(torch.tensor(st, dtype=dtype)); p.set_env_params(torch.tensor(size), torch.tensor(active))
    # what a caller of RobotArm.run_controller reads afterwards: the state, `converged` and the end-effector site (robot_arm.py:55-58,94)
    p.set_out_fields("SITE")
    flush = torch.empty(192 * 1024 * 1024, dtype=torch.uint8, device=dev)   # > 126 MB L2
    # pinned host staging for the end-to-end leg
    n_t = 8
    tgs = [sc.random_targets(n, seed=100 * rank + i) for i in range(n_t)]
    h_tg = [torch.tensor(t, dtype=dtype).pin_memory() for t, _ in tgs]
    h_gr = [torch.tensor(g).pin_memory() for _, g in tgs]
    d_tg = [t.to(dev) for t in h_tg]; d_gr = [g.to(dev) for g in h_gr]
    h_state = torch.empty_like(p.state, device="cpu").pin_memory(); h_conv = torch.empty(n, dtype=torch.uint8).pin_memory()
    if not wl["osc"]:
        p.forward(full_contacts=False); torch.cuda.synchronize()
        o_bias = p.data.qfrc_bias[0, :7].cpu().numpy().astype(np.float64)        # gravity load of the home pose: the kernel's own qfrc_bias
        ctrls = [torch.tensor(sc.halton_ctrl(min(n, 256), i, o_bias), dtype=dtype, device=dev).repeat((n + 255) // 256, 1)[:n].contiguous() for i in range(n_t)]
        h_ctrl = [c.cpu().pin_memory() for c in ctrls]

    def one_step(i, host_io=False):
        k = i % n_t
        if wl["osc"]:
            if host_io:
                d_tg[k].copy_(h_tg[k], non_blocking=True); d_gr[k].copy_(h_gr[k], non_blocking=True)
            p.run_controller(d_tg[k], d_gr[k], ticks, CONTROL_STEPS)
        else:
            if host_io:
                ctrls[k].copy_(h_ctrl[k], non_blocking=True)
            p.ctrl.copy_(ctrls[k]); p.step(steps_per_launch)
        if host_io:
            h_state.copy_(p.state, non_blocking=True); h_conv.copy_(p.converged, non_blocking=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(args.warmup):
        one_step(i)
    barrier()
    sampler = ClockSampler(local); sampler.start()
    launches0, order0 = p.launches, p.order_launches
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    wall0 = time.perf_counter()
    for i in range(args.steps):
        flush.fill_(i & 255)                      # evict L2 between timed iterations (not timed)
        ev[i][0].record(); one_step(args.warmup + i); ev[i][1].record()
    barrier()
    wall = time.perf_counter() - wall0
    kern_ms = [a.elapsed_time(b) for a, b in ev]
    launches = p.launches - launches0
    order_launches = p.order_launches - order0
    total_ms = torch.tensor([sum(kern_ms)], dtype=torch.float64, device=dev)
    # end-to-end leg: same call with HOST buffers (pinned), h2d of the action + d2h of state/converged inside the timed region
    e2e_steps = max(2, args.steps)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(e2e_steps):
        one_step(args.warmup + args.steps + i, host_io=True)
    e1.record(); barrier()
    e2e_ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    sampler.stop_flag = True; sampler.join(timeout=2)
    st_bits = p.status()
    status_bad = int((st_bits != 0).sum().item())
    flag_counts = [float(((st_bits & (1 << b)) != 0).sum().item()) for b in range(len(FLAG_NAMES))]
    stats = torch.tensor([float(p.converged.sum().item()), float(n), float(status_bad)] + flag_counts, dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX); dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)     # episode statistics: the only collective on this path
    if rank == 0:
        t_s = total_ms.item() * 1e-3
        total_env_steps = float(n) * world * steps_per_launch * args.steps
        value = total_env_steps / t_s
        e2e_value = float(n) * world * steps_per_launch * e2e_steps / (e2e_ms.item() * 1e-3)
        # roofline: algorithmic bytes per env-step, physics+OSC fused per tick: 258w / 5 steps ; physics-only 252w / step
        alg_b = (258 * w / CONTROL_STEPS) if wl["osc"] else 252 * w
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))); peak, which = float(peaks["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            peak, which = 6650.0, "fallback (B200_PROFILING.md)"
        per_launch_s = t_s / max(1, launches) if launches else t_s
        achieved = alg_b * n * steps_per_launch / per_launch_s / 1e9
        import ctypes as C
        from mujoco_robot_environments_b200 import ffi
        fma = C.c_double(0); ffi.lib().b2p_fma_peak(0 if wl["dtype"] == "f64" else 1, local, C.byref(fma))
        h2d = n * (13 * w + 1) if wl["osc"] else n * 8 * w
        line = {"metric": "env-steps/sec (physics+OSC)" if wl["osc"] else "env-steps/sec (physics only)", "value": value, "unit": "env-steps/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms.item() / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": wl["dtype"], "data": "synthetic",
                "config": {**workload_config(args, wl),
                           "l2": "flushed between timed iterations (192 MiB fill, untimed); the e2e leg runs back to back without the flush", "launch": p.launch_info(),
                           "derived_outputs": "stepping launches refresh the end-effector site group only (b2p_set_out_mask(B2P_OUT_SITE)); state, converged and status always", "wall_s_incl_flush": wall,
                           "episode_stats": {"converged_envs": stats[0].item(), "envs": stats[1].item(), "flagged_envs": stats[2].item(),
                                             "flags": {k: stats[3 + i].item() for i, k in enumerate(FLAG_NAMES)}}},
                "clocks": sampler.summary(), "gpu_launches": int(launches + order_launches) * world,
                "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(n * (128 * w + 1))},
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": measured_traffic(args.workload),
                             "peak_source": which, "algorithmic_bytes_per_env_step": alg_b,
                             "note": "path is latency / FP-issue bound, not HBM bound (SURVEY 7.3 H7); FP pipe figures below",
                             "fma_peak_tflops_measured": fma.value}}
        if world == 1 and not args.no_cpu:
            v, cores = cpu_baseline(args.cpu_seconds, wl["osc"])
            line["cpu_baseline"] = {"value": v, "unit": "env-steps/s", "cores": cores, "kind": "port",
                                    "sample": "%d processes x %.0f s, one env per process, same scene/target distribution (oracle C port, %s; MuJoCo not installable here)" % (cores, args.cpu_seconds, CPU_BUILD)}
        p.close()
        return line
    p.close()
    return None


def run_ours(args, wl):
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    line = measure(args, wl, args.workload, args.steps)
    if args.workload == "c3" and world == 1 and not args.no_secondary:
        # BASELINE configs[1] (fp64 parity mode, physics only) rides along on the default line so that a driver-run record carries it
        a2 = argparse.Namespace(**{**vars(args), "no_cpu": True})
        l2 = measure(a2, WORKLOADS["c2"], "c2", max(3, min(5, args.steps)))
        if line is not None and l2 is not None:
            line["secondary"] = {k: l2[k] for k in ("metric", "value", "unit", "dtype", "steps", "ms_per_step", "config", "e2e", "roofline", "gpu_launches")}
    if line is not None:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--envs", type=int, default=0, help="envs per GPU (default: workload's)")
    ap.add_argument("--ticks", type=int, default=20, help="control ticks per launch (x5 physics steps)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="c3 only: skip the c2 (fp64 parity mode) measurement carried as `secondary`")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, wl)
    elif args.workload == "render":
        run_render(args, wl)
    else:
        run_ours(args, wl)


if __name__ == "__main__":
    main()
