#!/bin/bash
# quick A/B: product build only (c3, c2) + stage profile; no test-suite.   gpurun -- 'bash tools/gpu_quick.sh TAG'
TAG=${1:-q}; O=gpurun_out; mkdir -p $O
bash tools/gpu_variants.sh $TAG base "${@:2}"
B2P_LIB=$PWD/mujoco_robot_environments_b200/csrc/variants/libb2p_prof.so timeout 300 python tools/stage_profile.py > $O/${TAG}_stage_c3.txt 2>&1; cat $O/${TAG}_stage_c3.txt
