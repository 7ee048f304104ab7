#!/bin/bash
# march length against lattice size after the per-plane source / probe masks (the fit of pick_fused_zc's constants)
python bench.py --pml 16 --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('pml16 value %.2f e2e %.2f frac %.3f ms %.4f' % (d['value'], d['e2e']['value'], d['roofline']['frac'], d['ms_per_step']))"
for lat in 256,256,256 320,320,320 384,384,384 448,448,448 512,512,512 640,640,640; do
for zc in 0 16 24 32 48 64; do
  FDTD_B200_TUNE=0,0,$zc python bench.py --lattice $lat --steps 20 --warmup 5 --passes 9 --no-cpu-baseline --no-slab-parity 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$lat zc $zc value %.2f frac %.3f ms %.4f' % (d['value'], d['roofline']['frac'], d['ms_per_step']))"
done; done
