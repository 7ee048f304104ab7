#!/bin/bash
python -m pytest tests/test_2d_gpu.py -m gpu -x -q 2>&1 | tail -4
for lat in 4096,4096,1 2048,2048,1 1000,1000,1 8192,8192,1 4000,3000,1; do
  python bench.py --config thin --lattice $lat --steps 50 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$lat value %.2f e2e %.2f frac %.3f us/step %.1f clocks %s' % (d['value'], d['e2e']['value'], d['roofline']['frac'], 1e3*d['ms_per_step'], d['clocks']['sm_mhz']))"
  FDTD_B200_STEP2D=0 python bench.py --config thin --lattice $lat --steps 50 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('   z-marching kernel: value %.2f frac %.3f' % (d['value'], d['roofline']['frac']))"
done
python bench.py --config thin --steps 50 --warmup 5 > gpurun_out/r02_case_thin_2d.json 2>/dev/null
ncu --set full --clock-control none --import-source on -k regex:k_step_2d -s 3 -c 1 -o gpurun_out/step2d_v3 -f python bench.py --config thin --steps 5 --warmup 3 --passes 5 --no-cpu-baseline > gpurun_out/ncu_2d.log 2>&1
