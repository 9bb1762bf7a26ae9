#!/bin/bash
# GPU pass for the ray-cast observations: parity tests, bench line, launch list, one full capture of the frame kernel.
TAG=${1:-r01}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 $O/${TAG}_pytest_gpu.log
timeout 600 python bench.py --workload render > $O/${TAG}_bench_render.json 2> $O/${TAG}_bench_render.err; echo "render rc=$?"; cat $O/${TAG}_bench_render.json; tail -5 $O/${TAG}_bench_render.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/${TAG}_launches_render.csv \
    python bench.py --workload render --steps 2 --warmup 3 --no-cpu > $O/${TAG}_ncu_launches_render.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_frame_kernel -s 3 -c 1 -f -o $O/${TAG}_render_frame \
    python bench.py --workload render --steps 1 --warmup 3 --no-cpu > $O/${TAG}_ncu_full_render.log 2>&1; echo "ncu full rc=$?"
ls -la $O | tail -8
