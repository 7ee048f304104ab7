# Mid-size lattices (BASELINE configs[1] is 256^3): planes per CTA march (zc) against the automatic choice.  Outputs under gpurun_out/.
for L in 256,256,256 384,384,384 512,512,512; do
  for ZC in 0 8 12 16 20 24 32 48 64; do
    FDTD_B200_TUNE="0,0,$ZC,-1,0" python bench.py --lattice $L --steps 200 --warmup 10 --no-cpu-baseline 2>> gpurun_out/zc.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$L zc=$ZC', round(d['value'],2), 'Gcells/s', round(d['ms_per_step'],4), 'ms')" >> gpurun_out/exp_zc_midsize_v8.txt
  done
done
cat gpurun_out/exp_zc_midsize_v8.txt
