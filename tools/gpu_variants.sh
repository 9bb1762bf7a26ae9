#!/bin/bash
# A/B bench of library variants (tools/build_variant.py) on one GPU box:  gpurun -- 'bash tools/gpu_variants.sh TAG name1 name2 ...'
# "base" = the product build.  Each variant: c3 (6 launches) and c2 (4 launches), device-timed value only.
TAG=$1; shift
O=gpurun_out; mkdir -p $O
for v in "$@"; do
  if [ "$v" = base ]; then unset B2P_LIB; else export B2P_LIB=$PWD/mujoco_robot_environments_b200/csrc/variants/libb2p_$v.so; fi
  timeout 300 python bench.py --no-cpu --no-secondary --steps 6 > $O/${TAG}_${v}_c3.json 2> $O/${TAG}_${v}_c3.err
  timeout 300 python bench.py --workload c2 --no-cpu --steps 4 > $O/${TAG}_${v}_c2.json 2> $O/${TAG}_${v}_c2.err
  python - <<PY
import json
for w in ("c3","c2"):
    try:
        l=json.load(open("$O/${TAG}_${v}_%s.json"%w)); print("$v", w, "%.3f M env-steps/s"%(l["value"]/1e6), "e2e %.3f"%(l["e2e"]["value"]/1e6), l["config"]["launch"], l["config"]["episode_stats"]["flags"])
    except Exception as e: print("$v", w, "FAILED", e)
PY
done
