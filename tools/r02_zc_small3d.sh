#!/bin/bash
# march length on small 3-D lattices (FDTD_B200_TUNE=0,0,zc; 0 = the library's choice)
for spec in ${SPECS:-"128,128,128 0" "128,128,128 4" "128,128,128 6" "128,128,128 7" "128,128,128 9" "128,128,128 16" "160,160,160 0" "160,160,160 8" "160,160,160 10" "160,160,160 14" "200,200,200 0" "200,200,200 10" "200,200,200 20" "200,200,200 34"}; do
  set -- $spec
  FDTD_B200_TUNE=0,0,$2 timeout 120 python bench.py --lattice $1 --steps 100 --warmup 10 --passes 9 --no-cpu-baseline --no-slab-parity 2>/dev/null | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('$1 zc=$2 value %.2f e2e %.2f us/step %.2f' % (d['value'], d['e2e']['value'], 1e3*d['ms_per_step']))
except Exception as e: print('$1 FAILED')"
done
