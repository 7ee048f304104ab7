# One GPU call that refreshes the evidence of the default (fused) path: the bench lines, the ncu launch list of the same
# command and one full capture of the step kernel (plain and PML instantiations).  Outputs under gpurun_out/.
set -x
python bench.py > gpurun_out/bench_n1_final.json 2> gpurun_out/bench_n1_final.err
python bench.py --pml 16 --steps 100 --no-cpu-baseline > gpurun_out/bench_n1_pml16_final.json 2>> gpurun_out/bench_n1_final.err
python bench.py --exact --steps 100 --no-cpu-baseline > gpurun_out/bench_n1_exact_final.json 2>> gpurun_out/bench_n1_final.err
python bench.py --lattice 1024,1024,128 --steps 100 --no-cpu-baseline > gpurun_out/bench_n1_1024x1024x128_final.json 2>> gpurun_out/bench_n1_final.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_final.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_step_fused -s 4 -c 1 -o gpurun_out/fused_final -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_step_fused -s 4 -c 1 -o gpurun_out/fused_pml_final -f python bench.py --pml 16 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full_pml.log 2>&1
