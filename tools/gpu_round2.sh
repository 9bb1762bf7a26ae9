#!/bin/bash
# One GPU pass of round 2 (run under gpurun from the repo root):  gpurun --timeout 1800 -- 'bash tools/gpu_round2.sh TAG [quick]'
TAG=${1:-r02}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $O/${TAG}_smi.txt 2>&1
timeout 120 python tools/replay_bad_state.py > $O/${TAG}_replay.log 2>&1; tail -4 $O/${TAG}_replay.log
[ "$3" = "notest" ] || timeout 1500 python -m pytest tests -m gpu -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest_gpu.log; tail -3 $O/${TAG}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/${TAG}_smoke.log
timeout 900 python bench.py > $O/${TAG}_bench_c3.json 2> $O/${TAG}_bench_c3.err; echo "c3 rc=$?"; cut -c1-400 $O/${TAG}_bench_c3.json
timeout 600 python bench.py --workload c2 --no-cpu > $O/${TAG}_bench_c2.json 2> $O/${TAG}_bench_c2.err; echo "c2 rc=$?"; cut -c1-400 $O/${TAG}_bench_c2.json
if [ "$2" != "quick" ]; then
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err; echo "ref rc=$?"; cut -c1-300 $O/${TAG}_bench_reference.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches_c3.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-secondary > $O/${TAG}_ncu_launches.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:b2p_step_kernel -s 3 -c 1 -f -o $O/${TAG}_step_c3 python bench.py --steps 1 --warmup 3 --ticks 4 --no-cpu --no-secondary > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:b2p_step_kernel -s 3 -c 1 -f -o $O/${TAG}_step_c2 python bench.py --workload c2 --steps 1 --warmup 3 --ticks 4 --no-cpu > $O/${TAG}_ncu_full_c2.log 2>&1; echo "ncu full c2 rc=$?"
fi
if [ -f mujoco_robot_environments_b200/csrc/variants/libb2p_prof.so ]; then
B2P_LIB=$PWD/mujoco_robot_environments_b200/csrc/variants/libb2p_prof.so timeout 300 python tools/stage_profile.py > $O/${TAG}_stage_c3.txt 2>&1; cat $O/${TAG}_stage_c3.txt
B2P_LIB=$PWD/mujoco_robot_environments_b200/csrc/variants/libb2p_prof.so timeout 300 python tools/stage_profile.py --workload c2 > $O/${TAG}_stage_c2.txt 2>&1
fi
ls $O | grep ${TAG}
