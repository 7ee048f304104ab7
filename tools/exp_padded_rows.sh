# Row padding: GPU test suite incl. tests/test_padded_rows_gpu.py, then the unaligned-lattice bench lines
# (padded rows → fused kernel) beside the same lattices with FDTD_B200_NO_ROW_PADDING=1 (one-thread-per-cell kernels),
# and the aligned headline as a regression check.  Outputs under gpurun_out/.
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_padded.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_padded.log
tail -5 gpurun_out/pytest_gpu_padded.log
python bench.py --steps 100 --no-cpu-baseline > gpurun_out/bench_n1_512_after_padding.json 2> gpurun_out/bench_pad.err
for L in 500,500,500 509,509,509 250,250,250 1000,1000,128; do
  python bench.py --lattice $L --steps 100 --no-cpu-baseline > gpurun_out/bench_n1_${L//,/x}_padded.json 2>> gpurun_out/bench_pad.err
  FDTD_B200_NO_ROW_PADDING=1 python bench.py --lattice $L --steps 20 --no-cpu-baseline > gpurun_out/bench_n1_${L//,/x}_unpadded.json 2>> gpurun_out/bench_pad.err
done
python bench.py --lattice 500,500,500 --pml 16 --steps 100 --no-cpu-baseline > gpurun_out/bench_n1_500x500x500_pml16_padded.json 2>> gpurun_out/bench_pad.err
tail -3 gpurun_out/bench_pad.err
cat gpurun_out/bench_n1_*padded.json gpurun_out/bench_n1_512_after_padding.json | cut -c1-400
