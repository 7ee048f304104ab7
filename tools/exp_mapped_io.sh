# One-step protocol through mapped pinned memory (forcing slot read by the kernels, probe ring written by them) against the
# staged copies (FDTD_B200_STAGED_IO=1): strict and read-lagging e2e at 512^3 and 256^3.
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for L in 512,512,512 256,256,256; do
  for rep in 1 2; do
  for S in 0 1; do
    FDTD_B200_STAGED_IO=$S python bench.py --lattice $L --steps 200 --warmup 10 --no-cpu-baseline 2>> gpurun_out/mapped.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$L staged=$S value', round(d['value'],2), 'e2e', round(d['e2e']['value'],2), 'lagging', round(d['e2e']['read_lagging_one_step']['value'],2))" >> gpurun_out/exp_mapped_io.txt
  done
  done
done
cat gpurun_out/exp_mapped_io.txt; tail -3 gpurun_out/mapped.err
