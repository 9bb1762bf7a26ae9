"""Summarise an ncu report of the step kernel into a small text file for profiles/.

  python tools/ncu_summary.py gpurun_out/step_c3.ncu-rep profiles/r01_step_c3_ncu.txt

Reads the report with `ncu -i ... --page raw --csv` (headline counters) and `--page source --csv --print-source
cuda,sass` (per-line instruction counts and stall samples, needs -lineinfo), and aggregates the per-line numbers by
the function of csrc/b2p_core.h they fall in.
"""
import collections
import csv
import io
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "mujoco_robot_environments_b200", "csrc")
RAW_KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread", "launch__grid_size",
            "launch__block_size", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
            "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
            "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sass__inst_executed_local_loads", "sass__inst_executed_local_stores",
            "smsp__average_warp_latency_per_inst_issued.ratio"]


def ncu_csv(rep, *args):
    out = subprocess.run(["ncu", "-i", rep, "--csv", *args], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def functions():
    """{file: [(first line, function name), ...]} for every header of csrc/."""
    out = {}
    for f in os.listdir(CSRC):
        if not f.endswith((".h", ".cuh")):
            continue
        funcs = []
        for i, l in enumerate(open(os.path.join(CSRC, f)).read().split("\n"), 1):
            if "B2P_HD" in l and (l.startswith("B2P_HD") or l.startswith("template")):
                m = re.search(r"(\w+)\(", l.split("B2P_HD", 1)[1])
                if m:
                    funcs.append((i, m.group(1)))
        out[f] = funcs
    return out


def main():
    rep, dst = sys.argv[1], sys.argv[2]
    lines = ["ncu summary of %s" % os.path.basename(rep), ""]
    raw = ncu_csv(rep, "--page", "raw")
    hdr, units, vals = raw[0], raw[1], raw[2]
    kname = vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
    lines.append("kernel: %s" % kname)
    for h, u, v in zip(hdr, units, vals):
        if h in RAW_KEYS or h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio"):
            lines.append("  %-80s %s %s" % (h, v, u))
    src = ncu_csv(rep, "--page", "source", "--print-source", "cuda,sass")
    funcs = functions()

    def fn(f, line):
        name = f
        for i, n in funcs.get(f, []):
            if i <= line:
                name = n
            else:
                break
        return name
    cur, h, agg = None, None, {}
    for r in src:
        if len(r) >= 2 and r[0] == "File Path":
            cur = r[1].split("/")[-1]
        elif len(r) > 3 and r[0] == "Line No":
            h = r
        elif h and len(r) == len(h) and r[0].isdigit() and r[2] == "-":
            agg[(cur, int(r[0]))] = (int(r[h.index("Instructions Executed")] or 0), int(r[h.index("# Samples")] or 0), r[1].strip()[:100])
    ti, ts = sum(v[0] for v in agg.values()) or 1, sum(v[1] for v in agg.values()) or 1
    fa, fs = collections.Counter(), collections.Counter()
    for (f, l), (ni, ns, _) in agg.items():
        k = fn(f, l)
        fa[k] += ni
        fs[k] += ns
    lines += ["", "warp instructions executed: %d   stall samples: %d" % (ti, ts), "", "by function (share of instructions / share of samples):"]
    for k, v in fa.most_common(28):
        lines.append("  %-26s %5.1f%%  %5.1f%%" % (k, 100.0 * v / ti, 100.0 * fs[k] / ts))
    lines += ["", "top source lines by samples:"]
    for (f, l), (ni, ns, s) in sorted(agg.items(), key=lambda x: -x[1][1])[:25]:
        lines.append("  %s:%d  inst %.1f%%  samples %.1f%%  | %s" % (f, l, 100.0 * ni / ti, 100.0 * ns / ts, s[:80]))
    with open(dst, "w") as f:
        f.write("\n".join(lines) + "\n")
    print("\n".join(lines[:60]))


if __name__ == "__main__":
    main()
