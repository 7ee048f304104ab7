# Where the fused kernel requests its first planes: 0 at kernel start (lane that initialises the barriers) | 1 after the first
# CTA barrier | 2 after the forcing / probe scans (the earlier builds).  Same box, alternating.
for L in 512,512,512 256,256,256 384,384,384 1024,1024,128; do
  for rep in 1 2 3; do
  for A in 0 1 2; do
    FDTD_B200_ISSUE_MODE=$A python bench.py --lattice $L --steps 200 --warmup 10 --no-cpu-baseline 2>> gpurun_out/early.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$L issue_mode=$A', round(d['value'],2), 'Gcells/s', round(d['ms_per_step'],4), 'ms')" >> gpurun_out/exp_early_issue3.txt
  done
  done
done
cat gpurun_out/exp_early_issue3.txt; tail -3 gpurun_out/early.err
