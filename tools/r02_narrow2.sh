#!/bin/bash
timeout 600 python -m pytest tests/test_narrow_tiles_gpu.py -m gpu -x -q 2>&1 | tail -3
for lat in 512,512,512 320,320,320 448,448,448 192,192,192; do
  FDTD_B200_NARROW_TILES=1 timeout 300 python bench.py --lattice $lat --steps 20 --warmup 5 --passes 9 --no-cpu-baseline --no-slab-parity 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$lat narrow=1 value %.2f e2e %.2f frac %.3f ms %.4f' % (d['value'], d['e2e']['value'], d['roofline']['frac'], d['ms_per_step']))"
done
FDTD_B200_NARROW_TILES=1 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:k_step_fused -s 8 -c 1 --csv --log-file gpurun_out/narrow_512.csv python bench.py --steps 3 --warmup 3 --passes 5 --no-cpu-baseline --no-slab-parity > /dev/null 2>&1
python - gpurun_out/narrow_512.csv <<'PY'
import csv,sys
rows=[r for r in csv.reader(open(sys.argv[1])) if r and not r[0].startswith('==')]
h=rows[0]
print("   512 narrow:", ", ".join(f"{r[h.index('Metric Name')]} {r[h.index('Metric Value')]} {r[h.index('Metric Unit')]}" for r in rows[1:]))
PY
