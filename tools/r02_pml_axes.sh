#!/bin/bash
# which slab axis costs what: bench value + DRAM bytes + warp instructions of the PML instantiation per axis subset
for ax in ${AXES:-x y z xyz}; do
  python bench.py --pml 16 --pml-axes $ax --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('axes $ax value %.2f frac %.3f ms %.4f psi_bytes %.3f GB' % (d['value'], d['roofline']['frac'], d['ms_per_step'], d['impl_detail']['cpml_psi_bytes_per_step']/1e9))"
  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,gpu__time_duration.sum --clock-control none -k regex:k_step_fused -s 8 -c 1 --csv --log-file gpurun_out/ax_$ax.csv \
      python bench.py --pml 16 --pml-axes $ax --steps 3 --warmup 3 --passes 5 --no-cpu-baseline > /dev/null 2>&1
  python - gpurun_out/ax_$ax.csv <<'PY'
import csv,sys
rows=[r for r in csv.reader(open(sys.argv[1])) if r and not r[0].startswith('==')]
h=rows[0]
print("   ", ", ".join(f"{r[h.index('Metric Name')]} {r[h.index('Metric Value')]} {r[h.index('Metric Unit')]}" for r in rows[1:]))
PY
done
