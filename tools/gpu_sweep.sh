#!/bin/bash
# BASELINE configs[4] on one GPU: c3 (default scene + OSC, fp32) at 1k..64k scenes per GPU; plus the c2 flag histogram.
TAG=${1:-r01}
O=gpurun_out
mkdir -p $O
: > $O/${TAG}_sweep_c3.jsonl
for E in 1024 2048 4096 8192 16384 32768 65536; do
  timeout 600 python bench.py --envs $E --steps 5 --no-cpu >> $O/${TAG}_sweep_c3.jsonl 2>> $O/${TAG}_sweep.err; echo "envs $E rc=$?"
done
timeout 600 python bench.py --workload c2 --no-cpu > $O/${TAG}_bench_c2.json 2>> $O/${TAG}_sweep.err; echo "c2 rc=$?"
python - $TAG <<'P'
import json,sys
for l in open("gpurun_out/%s_sweep_c3.jsonl" % sys.argv[1] if len(sys.argv)>1 else "gpurun_out/r01_sweep_c3.jsonl"):
    d=json.loads(l); print(d["config"]["envs_per_gpu"], round(d["value"]), round(d["e2e"]["value"]), d["config"]["episode_stats"]["flags"])
P
python -c "
import json; d=json.load(open('$O/${TAG}_bench_c2.json')); print('c2', round(d['value']), d['config']['episode_stats'])"
