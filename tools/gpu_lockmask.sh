#!/bin/bash
# lockstep-mask sweep on the product build (B2P_LOCKSTEP bit mask of the CTA-wide stage barriers, csrc/b2p_math.h LK_*)
TAG=$1; shift; O=gpurun_out; mkdir -p $O
for m in "$@"; do
  B2P_LOCKSTEP=$m timeout 300 python bench.py --no-cpu --no-secondary --steps 8 > $O/${TAG}_lk${m}_c3.json 2>/dev/null
  python -c "import json; l=json.load(open('$O/${TAG}_lk${m}_c3.json')); print('mask $m c3 %.3f M  e2e %.3f' % (l['value']/1e6, l['e2e']['value']/1e6))"
done
