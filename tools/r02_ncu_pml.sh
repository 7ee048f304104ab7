#!/bin/bash
# CPML instantiation of the fused kernel: plain bench line, then one full capture with source counters; $1 = tag, AX = axes (default xyz)
AX=${AX:-xyz}
python bench.py --pml 16 --pml-axes $AX --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/pml_plain_${1:-a}.json 2> gpurun_out/pml_plain.err || exit 1
python -c "
import json; d=json.load(open('gpurun_out/pml_plain_${1:-a}.json')); print('pml16 $AX value %.2f e2e %.2f frac %.3f ms %.4f' % (d['value'], d['e2e']['value'], d['roofline']['frac'], d['ms_per_step']))"
ncu --set full --clock-control none --import-source on -k regex:k_step_fused -s 8 -c 1 -o gpurun_out/r02_pml_${1:-a} -f \
    python bench.py --pml 16 --pml-axes $AX --steps 3 --warmup 3 --passes 5 --no-cpu-baseline > gpurun_out/ncu_pml.log 2>&1
ls -la gpurun_out/r02_pml_${1:-a}.ncu-rep
