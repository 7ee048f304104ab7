#!/bin/bash
# march length on the 64-cell tile, 20 steps per pass (short enough to stay out of the power cap): FDTD_B200_TUNE=0,0,zc forces it
for spec in "300,300,300 0" "300,300,300 15" "300,300,300 28" "320,320,320 0" "448,448,448 0" "192,192,192 0" "192,192,192 28"; do
  set -- $spec
  FDTD_B200_TUNE=0,0,$2 timeout 120 python bench.py --lattice $1 --steps 20 --warmup 5 --passes 9 --no-cpu-baseline --no-slab-parity 2>/dev/null | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('$1 zc $2 value %.2f e2e %.2f frac %.3f ms %.4f clocks %s %s' % (d['value'], d['e2e']['value'], d['roofline']['frac'], d['ms_per_step'], d['clocks']['sm_mhz'], d['clocks']['reasons']))
except Exception as e: print('$1 FAILED')"
done
