# A/B of the tensor maps' L2 promotion (FDTD_B200_TUNE field 7: 0 = 256 B (default), 1 none, 2 64 B, 3 128 B; measured when 0 was 128 B and 3 was 256 B), alternating runs
for rep in 1 2; do for pr in 0 1 2 3; do
  FDTD_B200_TUNE=0,0,0,-1,0,0,$pr python bench.py --steps 100 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('RES promo', $pr, round(d['value'],2), round(d['roofline']['kernel_avg_ms'],4))"
done; done
for pr in 0 1 3; do
  FDTD_B200_TUNE=0,0,0,-1,0,0,$pr python bench.py --lattice 1024,1024,128 --steps 100 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('RES1024 promo', $pr, round(d['value'],2), round(d['roofline']['kernel_avg_ms'],4))"
done
