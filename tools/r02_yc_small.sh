#!/bin/bash
# rows per march on small 2-D lattices
for spec in ${SPECS:-"1400,1400,1 8" "1400,1400,1 4" "1700,1700,1 8" "1700,1700,1 6" "1700,1700,1 4" "800,800,1 2" "800,800,1 3" "800,800,1 4"}; do
  set -- $spec
  FDTD_B200_STEP2D_YC=$2 timeout 120 python bench.py --config thin --lattice $1 --steps 100 --warmup 10 --passes 9 --no-cpu-baseline --no-slab-parity 2>/dev/null | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('$1 yc=$2 value %.2f e2e %.2f us/step %.2f' % (d['value'], d['e2e']['value'], 1e3*d['ms_per_step']))
except Exception as e: print('$1 FAILED')"
done
