#!/bin/bash
timeout 300 python -m pytest tests/test_2d_gpu.py tests/test_fused_gpu.py -m gpu -x -q --timeout 60 2>&1 | tail -2
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1 | cut -c1-80
for lat in 600,600,1 800,800,1 1000,1000,1 1400,1400,1 1700,1700,1 2048,2048,1 4096,4096,1; do
  timeout 120 python bench.py --config thin --lattice $lat --steps 50 --warmup 10 --passes 9 --no-cpu-baseline --no-slab-parity 2>/dev/null | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('$lat value %.2f e2e %.2f us/step %.2f' % (d['value'], d['e2e']['value'], 1e3*d['ms_per_step']))
except Exception as e: print('$lat FAILED')"
done
