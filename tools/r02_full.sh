#!/bin/bash
# whole GPU suite + smoke + the committed bench lines of the round (1 GPU)
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_n1_final.json 2>gpurun_out/final.err
python bench.py --config 2 --steps 50 --warmup 5 > gpurun_out/r02_case_2_masks.json 2>>gpurun_out/final.err
python bench.py --lattice 256,256,256 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02_bench_n1_256_masks.json 2>>gpurun_out/final.err
python bench.py --lattice 384,384,384 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02_bench_n1_384_masks.json 2>>gpurun_out/final.err
python bench.py --pml 16 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02_bench_n1_pml16.json 2>>gpurun_out/final.err
python bench.py --config thin --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r02_case_thin_2d.json 2>>gpurun_out/final.err
python bench.py --config app --steps 200 --warmup 5 --no-cpu-baseline > gpurun_out/r02_case_app_final.json 2>>gpurun_out/final.err
for f in gpurun_out/r02_bench_n1_final.json gpurun_out/r02_case_2_masks.json gpurun_out/r02_bench_n1_256_masks.json gpurun_out/r02_bench_n1_384_masks.json gpurun_out/r02_bench_n1_pml16.json gpurun_out/r02_case_thin_2d.json gpurun_out/r02_case_app_final.json; do python -c "
import json,sys
d=json.load(open('$f')); print('$f', d['config']['lattice'], 'value %.2f e2e %.2f frac %.3f ms %.4f' % (d['value'], d['e2e']['value'], d['roofline']['frac'], d['ms_per_step']))"; done
tail -3 gpurun_out/final.err
