#!/bin/bash
# lockstep-group sweep on the product build (B2P_LOCKGROUP = warps per lockstep group, 0 = the whole CTA; csrc/b2p_kernel.cuh)
#   gpurun -- 'bash tools/gpu_lockgroup.sh TAG c3|c2 groups...'
TAG=$1; WL=$2; shift; shift; O=gpurun_out; mkdir -p $O
for m in "$@"; do
  B2P_LOCKGROUP=$m timeout 300 python bench.py --workload $WL --no-cpu --no-secondary --steps 8 > $O/${TAG}_lg${m}_${WL}.json 2>/dev/null
  python -c "import json; l=json.load(open('$O/${TAG}_lg${m}_${WL}.json')); print('group $m $WL %.3f M  e2e %.3f' % (l['value']/1e6, l['e2e']['value']/1e6))"
done
