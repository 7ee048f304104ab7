#!/bin/bash
# whole GPU suite + smoke + the committed bench lines of the round (1 GPU); every step under its own timeout
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
b() { out=$1; shift; timeout 240 python bench.py "$@" > gpurun_out/$out.json 2>>gpurun_out/final.err; }
: > gpurun_out/final.err
b r02_bench_n1_final --steps 20 --warmup 5
b r02_case_2_masks --config 2 --steps 50 --warmup 5
b r02_bench_n1_256_masks --lattice 256,256,256 --steps 20 --warmup 5 --no-cpu-baseline
b r02_bench_n1_384_masks --lattice 384,384,384 --steps 20 --warmup 5 --no-cpu-baseline
b r02_bench_n1_pml16 --pml 16 --steps 20 --warmup 5 --no-cpu-baseline
b r02_case_thin_2d --config thin --steps 50 --warmup 5 --no-cpu-baseline
b r02_case_app_final --config app --steps 200 --warmup 5 --no-cpu-baseline
b r02_case_1_final --config 1 --steps 200 --warmup 5 --no-cpu-baseline
for f in r02_bench_n1_final r02_case_2_masks r02_bench_n1_256_masks r02_bench_n1_384_masks r02_bench_n1_pml16 r02_case_thin_2d r02_case_app_final r02_case_1_final; do python -c "
import json,sys
try:
    d=json.load(open('gpurun_out/$f.json')); print('$f', d['config']['lattice'], 'value %.2f e2e %.2f frac %.3f ms %.4f' % (d['value'], d['e2e']['value'], d['roofline']['frac'], d['ms_per_step']))
except Exception as e: print('$f FAILED', e)"; done
tail -3 gpurun_out/final.err
