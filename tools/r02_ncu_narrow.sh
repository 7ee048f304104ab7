#!/bin/bash
FDTD_B200_NARROW_TILES=1 python bench.py --steps 3 --warmup 3 --passes 5 --no-cpu-baseline --no-slab-parity > /dev/null 2>&1 || exit 1
FDTD_B200_NARROW_TILES=1 ncu --set full --clock-control none --import-source on -k regex:k_step_fused -s 8 -c 1 -o gpurun_out/r02_narrow512 -f \
    python bench.py --steps 3 --warmup 3 --passes 5 --no-cpu-baseline --no-slab-parity > gpurun_out/ncu_narrow.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_step_fused -s 8 -c 1 -o gpurun_out/r02_wide512 -f \
    python bench.py --steps 3 --warmup 3 --passes 5 --no-cpu-baseline --no-slab-parity > gpurun_out/ncu_wide.log 2>&1
ls -la gpurun_out/*.ncu-rep
