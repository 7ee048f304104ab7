# The automatic planes-per-march choice (pick_fused_zc) on a range of lattices, plain and with a 16-cell CPML.  Outputs under gpurun_out/.
python -m pytest tests/test_fused_gpu.py tests/test_padded_rows_gpu.py tests/test_pml_gpu.py -m gpu -x -q 2>&1 | tail -2
for L in 128,128,128 250,250,250 256,256,256 300,300,300 384,384,384 512,512,512 640,640,640 1000,1000,128 1024,1024,128; do
  for P in 0 16; do
    python bench.py --lattice $L --pml $P --steps 200 --warmup 10 --no-cpu-baseline 2>> gpurun_out/zc_auto.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$L pml=$P', round(d['value'],2), 'Gcells/s', round(d['ms_per_step'],4), 'ms', 'e2e', round(d['e2e']['value'],2), 'frac', round(d['roofline']['frac'],3))" >> gpurun_out/exp_zc_auto.txt
  done
done
cat gpurun_out/exp_zc_auto.txt; tail -3 gpurun_out/zc_auto.err
