#!/bin/bash
# round-2 check on a 2-GPU box: multi-GPU parity tests, N=2 bench (chunk-aligned vs one-plane boundary launches), N=1 bench + reference arm
set -u
O=gpurun_out
mkdir -p $O
python -m pytest tests/test_multigpu.py tests/test_fused_gpu.py tests/test_pml_gpu.py -m gpu -x -q 2>&1 | tail -15 > $O/r02_pytest_n2.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29600"
$TR bench.py --gpus 2 --steps 100 --warmup 10 > $O/r02_bench_n2_chunks.json 2> $O/r02_bench_n2_chunks.err
FDTD_B200_P2P_PLANE_SPLIT=1 $TR bench.py --gpus 2 --steps 100 --warmup 10 --no-slab-parity > $O/r02_bench_n2_planesplit.json 2> $O/r02_bench_n2_planesplit.err
# a middle rank has two neighbours: emulate its slab shape on 2 GPUs is not possible; N=2 lattices of 1024x1024x256 show the plane shape
$TR bench.py --gpus 2 --steps 100 --warmup 10 --lattice 1024,1024,256 --no-slab-parity > $O/r02_bench_n2_1024x1024x256_chunks.json 2>> $O/r02_bench_n2_chunks.err
FDTD_B200_P2P_PLANE_SPLIT=1 $TR bench.py --gpus 2 --steps 100 --warmup 10 --lattice 1024,1024,256 --no-slab-parity > $O/r02_bench_n2_1024x1024x256_planesplit.json 2>> $O/r02_bench_n2_planesplit.err
python bench.py --steps 20 --warmup 5 > $O/r02_bench_n1_k20.json 2> $O/r02_bench_n1_k20.err
python bench.py --impl reference --steps 20 --warmup 5 > $O/r02_bench_reference_n1.json 2> $O/r02_bench_reference_n1.err
$TR bench.py --impl reference --gpus 2 --steps 20 --warmup 5 > $O/r02_bench_reference_n2.json 2> $O/r02_bench_reference_n2.err
tail -3 $O/r02_pytest_n2.log
for f in $O/r02_bench_n2_chunks.json $O/r02_bench_n2_planesplit.json $O/r02_bench_n2_1024x1024x256_chunks.json $O/r02_bench_n2_1024x1024x256_planesplit.json $O/r02_bench_n1_k20.json $O/r02_bench_reference_n1.json $O/r02_bench_reference_n2.json; do echo $f; cut -c1-200 $f; done
