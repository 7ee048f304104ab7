#!/bin/bash
# round-2 check on a 1-GPU box: GPU suite, smoke, N=1 bench (driver's K/W) + reference arm on the whole lattice
set -u
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > $O/r02_pytest_n1.log
python __graft_entry__.py smoke > $O/r02_smoke.log 2>&1
python bench.py --steps 20 --warmup 5 > $O/r02_bench_n1_k20.json 2> $O/r02_bench_n1_k20.err
python bench.py --impl reference --steps 20 --warmup 5 > $O/r02_bench_reference_n1.json 2> $O/r02_bench_reference_n1.err
nproc > $O/r02_host.txt; free -g >> $O/r02_host.txt; lscpu | head -20 >> $O/r02_host.txt
tail -3 $O/r02_pytest_n1.log; tail -2 $O/r02_smoke.log
for f in $O/r02_bench_n1_k20.json $O/r02_bench_reference_n1.json; do echo $f; cut -c1-250 $f; done
cat $O/*.err | tail -20
