#!/bin/bash
# same-box A/B of library variants, interleaved twice (run-to-run spread on one box is visible):  gpurun -- 'bash tools/gpu_ab.sh TAG v1 v2 ...'
TAG=$1; shift
O=gpurun_out; mkdir -p $O
for rep in 1 2; do
for v in "$@"; do
  export B2P_LIB=$PWD/mujoco_robot_environments_b200/csrc/variants/libb2p_$v.so
  timeout 300 python bench.py --no-cpu --no-secondary --steps 8 > $O/${TAG}_${v}_${rep}_c3.json 2>/dev/null
  python -c "import json; l=json.load(open('$O/${TAG}_${v}_${rep}_c3.json')); print('$v rep$rep c3 %.3f M  e2e %.3f' % (l['value']/1e6, l['e2e']['value']/1e6))"
done
done
