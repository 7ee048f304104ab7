#!/bin/bash
python -m pytest tests/test_fused_gpu.py tests/test_parity_gpu.py tests/test_pml_gpu.py -m gpu -x -q 2>&1 | tail -3
for spec in "256,256,256 0" "256,256,256 16" "256,256,256 32" "256,256,256 64" "384,384,384 0" "512,512,512 0"; do
  set -- $spec
  FDTD_B200_TUNE=0,0,$2 python bench.py --lattice $1 --steps 20 --warmup 5 --no-cpu-baseline --no-slab-parity 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$1 zc $2 value %.2f e2e %.2f frac %.3f ms %.4f' % (d['value'], d['e2e']['value'], d['roofline']['frac'], d['ms_per_step']))"
done
python bench.py --config 2 --steps 50 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('config 2 value %.2f e2e %.2f frac %.3f ms %.4f' % (d['value'], d['e2e']['value'], d['roofline']['frac'], d['ms_per_step']))"
python bench.py --pml 16 --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('pml16 value %.2f e2e %.2f frac %.3f ms %.4f' % (d['value'], d['e2e']['value'], d['roofline']['frac'], d['ms_per_step']))"
