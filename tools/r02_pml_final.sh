#!/bin/bash
python -m pytest tests/test_pml_gpu.py tests/test_fused_gpu.py -m gpu -x -q 2>&1 | tail -3
for lat in 512,512,512 256,256,256 640,640,640 500,500,500; do
python bench.py --pml 16 --lattice $lat --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('pml16 $lat value %.2f e2e %.2f frac %.3f ms %.4f' % (d['value'], d['e2e']['value'], d['roofline']['frac'], d['ms_per_step']))"
done
python bench.py --pml 16 --steps 20 --warmup 5 > gpurun_out/r02_bench_n1_pml16.json 2>/dev/null
AX=xyz bash tools/r02_ncu_pml.sh final
