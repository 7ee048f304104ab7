#!/bin/bash
timeout 600 python -m pytest tests/test_narrow_tiles_gpu.py tests/test_fused_gpu.py -m gpu -x -q 2>&1 | tail -6
for lat in 320,320,320 448,448,448 300,300,300 576,576,576 192,192,192 512,512,512; do
for nt in 0 1; do
  FDTD_B200_NARROW_TILES=$nt timeout 300 python bench.py --lattice $lat --steps 20 --warmup 5 --passes 9 --no-cpu-baseline --no-slab-parity 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$lat narrow=$nt value %.2f e2e %.2f frac %.3f ms %.4f' % (d['value'], d['e2e']['value'], d['roofline']['frac'], d['ms_per_step']))"
done; done
