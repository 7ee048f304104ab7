#!/bin/bash
# drift theory: shorter marches should cut the excess DRAM reads of the x-slab case
for ax in ${AXES:-x xyz}; do
for zc in ${ZCS:-8 16 32 64}; do
  export FDTD_B200_TUNE=0,0,$zc
  python bench.py --pml 16 --pml-axes $ax --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('axes $ax zc $zc value %.2f frac %.3f ms %.4f' % (d['value'], d['roofline']['frac'], d['ms_per_step']))"
  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,gpu__time_duration.sum --clock-control none -k regex:k_step_fused -s 8 -c 1 --csv --log-file gpurun_out/zc_$ax_$zc.csv \
      python bench.py --pml 16 --pml-axes $ax --steps 3 --warmup 3 --passes 5 --no-cpu-baseline > /dev/null 2>&1
  python - gpurun_out/zc_$ax_$zc.csv <<'PY'
import csv,sys
rows=[r for r in csv.reader(open(sys.argv[1])) if r and not r[0].startswith('==')]
h=rows[0]
print("   ", ", ".join(f"{r[h.index('Metric Name')]} {r[h.index('Metric Value')]} {r[h.index('Metric Unit')]}" for r in rows[1:]))
PY
done; done
unset FDTD_B200_TUNE
for zc in 16 32 64; do
  FDTD_B200_TUNE=0,0,$zc python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('plain zc $zc value %.2f frac %.3f ms %.4f' % (d['value'], d['roofline']['frac'], d['ms_per_step']))"
done
