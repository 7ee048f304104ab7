#!/bin/bash
# whole GPU suite + smoke + the committed bench lines of the round (1 GPU)
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -12
python bench.py --config thin --steps 50 --warmup 5 > gpurun_out/r02_case_thin_2d.json 2>gpurun_out/thin.err
python bench.py --config thin --lattice 8192,8192,1 --steps 20 --warmup 5 > gpurun_out/r02_case_thin_8192.json 2>>gpurun_out/thin.err
python bench.py --config thin --lattice 2048,2048,1 --steps 50 --warmup 5 > gpurun_out/r02_case_thin_2048.json 2>>gpurun_out/thin.err
python bench.py --config thin --lattice 1000,1000,1 --steps 50 --warmup 5 > gpurun_out/r02_case_thin_1000.json 2>>gpurun_out/thin.err
for f in gpurun_out/r02_case_thin_*.json; do python -c "
import json,sys
d=json.load(open('$f')); print('$f', d['config']['lattice'], 'value %.2f e2e %.2f frac %.3f traffic %s cpu %s' % (d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['traffic'], d.get('cpu_baseline',{}).get('value')))"; done
