#!/bin/bash
# round-2 ncu evidence (each capture after a plain run of the same command exited 0); $1 = part (a | b): gpurun_out/ returns <= 64 MiB per call
set -x
if [ "${1:-a}" = a ]; then
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_plain.json 2> gpurun_out/ncu_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_ncu_launches_bench.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_step_fused -s 8 -c 1 -o gpurun_out/r02_fused -f \
    python bench.py --steps 3 --warmup 3 --passes 5 --no-cpu-baseline > gpurun_out/ncu_fused.log 2>&1
bash tools/r02_ncu_pml.sh a
else
python bench.py --config app --steps 200 --warmup 5 --no-cpu-baseline > gpurun_out/ncu_app_plain.json 2> gpurun_out/ncu_plain_b.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_steps_cluster -s 2 -c 1 -o gpurun_out/r02_cluster -f \
    python bench.py --config app --steps 200 --warmup 5 --passes 5 --no-cpu-baseline > gpurun_out/ncu_cluster.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_steps_resident -s 2 -c 1 -o gpurun_out/r02_resident -f \
    python bench.py --config 1 --steps 200 --warmup 5 --passes 5 --no-cpu-baseline > gpurun_out/ncu_resident.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02_ncu_launches_app.csv \
    python bench.py --config app --steps 200 --warmup 5 --passes 5 --no-cpu-baseline > gpurun_out/ncu_list_app.log 2>&1
fi
ls -la gpurun_out/*.ncu-rep
