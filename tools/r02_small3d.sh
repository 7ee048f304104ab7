#!/bin/bash
# small 3-D lattices above the on-chip budget: which path, how fast
for lat in ${LATS:-80,80,80 96,96,96 112,112,112 128,128,128 160,160,160 200,200,200}; do
  timeout 120 python bench.py --lattice $lat --steps 100 --warmup 10 --passes 9 --no-cpu-baseline --no-slab-parity 2>/dev/null | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('$lat value %.2f e2e %.2f frac %.3f us/step %.2f | %s' % (d['value'], d['e2e']['value'], d['roofline']['frac'], 1e3*d['ms_per_step'], d['impl_detail']['step_path'][:150]))
except Exception as e: print('$lat FAILED')"
done
