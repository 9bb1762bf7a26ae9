#!/bin/bash
# GPU pass: full GPU test-suite, variant A/B, stage profile
TAG=${1:-r02b}; O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest_gpu.log; tail -5 $O/${TAG}_pytest_gpu.log
bash tools/gpu_variants.sh $TAG base w13 w14
B2P_LOCKSTEP=511 timeout 300 python bench.py --no-cpu --no-secondary --steps 6 > $O/${TAG}_lk511_c3.json 2>/dev/null; python -c "import json; l=json.load(open('$O/${TAG}_lk511_c3.json')); print('lockstep 511 c3', l['value']/1e6)"
B2P_LIB=$PWD/mujoco_robot_environments_b200/csrc/variants/libb2p_prof.so timeout 300 python tools/stage_profile.py > $O/${TAG}_stage_c3.txt 2>&1; cat $O/${TAG}_stage_c3.txt
B2P_LIB=$PWD/mujoco_robot_environments_b200/csrc/variants/libb2p_prof.so timeout 300 python tools/stage_profile.py --workload c2 > $O/${TAG}_stage_c2.txt 2>&1; cat $O/${TAG}_stage_c2.txt
