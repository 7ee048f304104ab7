# A/B of the fused kernel's cache hints (FDTD_B200_TUNE fields 7 = stores, 8 = TMA loads), alternating runs on one box
for rep in 1 2; do
for sh in 0 1 2; do for lh in 0 1 2; do
  FDTD_B200_TUNE=0,0,0,-1,0,0,$sh,$lh python bench.py --steps 100 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('RES store', $sh, 'load', $lh, round(d['value'],2), round(d['roofline']['kernel_avg_ms'],4))"
done; done; done
for sh in 0 1; do for lh in 0 1; do
  FDTD_B200_TUNE=0,0,0,-1,0,0,$sh,$lh python bench.py --lattice 1024,1024,128 --steps 100 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('RES1024 store', $sh, 'load', $lh, round(d['value'],2), round(d['roofline']['kernel_avg_ms'],4))"
done; done
