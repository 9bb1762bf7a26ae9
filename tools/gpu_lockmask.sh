#!/bin/bash
# lockstep-mask sweep on the product build (B2P_LOCKSTEP bit mask of the CTA-wide stage barriers, csrc/b2p_math.h LK_*)
#   gpurun -- 'bash tools/gpu_lockmask.sh TAG c3|c2 mask...'
TAG=$1; WL=$2; shift; shift; O=gpurun_out; mkdir -p $O
for m in "$@"; do
  B2P_LOCKSTEP=$m timeout 300 python bench.py --workload $WL --no-cpu --no-secondary --steps 8 > $O/${TAG}_lk${m}_${WL}.json 2>/dev/null
  python -c "import json; l=json.load(open('$O/${TAG}_lk${m}_${WL}.json')); print('mask $m $WL %.3f M  e2e %.3f' % (l['value']/1e6, l['e2e']['value']/1e6))"
done
