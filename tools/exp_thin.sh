# Thin and two-dimensional lattices (the application's own test lattice is 100x100x1): fused step vs two-kernel path.
for L in 4096,4096,1 2048,2048,4 1024,1024,16 4000,4000,1 100,100,1; do
  for F in "" "--no-fused"; do
    python bench.py --lattice $L --steps 200 --warmup 10 --no-cpu-baseline $F 2>> gpurun_out/thin.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$L $F', round(d['value'],2), 'Gcells/s', round(d['ms_per_step'],4), 'ms', d['config']['step_kernel'][:20], 'e2e', round(d['e2e']['value'],2))" >> gpurun_out/exp_thin.txt
  done
done
cat gpurun_out/exp_thin.txt; tail -3 gpurun_out/thin.err
