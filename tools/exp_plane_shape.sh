set -x
for lat in 512,512,512 1024,1024,128 1024,512,256 2048,512,128 512,2048,128 2048,2048,32; do
  for zc in 0 32 128; do
    FDTD_B200_TUNE=0,0,$zc python bench.py --lattice $lat --steps 100 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('RES', '$lat', $zc, round(d['value'],2), round(d['ms_per_step'],4))"
  done
done
ncu --set full --clock-control none --import-source on -k regex:k_step_fused -s 4 -c 1 -o gpurun_out/fused_1024sq -f python bench.py --lattice 1024,1024,128 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_1024sq.log 2>&1
