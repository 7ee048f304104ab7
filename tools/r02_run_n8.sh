#!/bin/bash
# round-2 scaling evidence on an 8-GPU box: N=8 and N=4 weak-scaling lines, chunk-aligned boundary launches (default) against the
# round-1 one-plane boundary launches (FDTD_B200_P2P_PLANE_SPLIT=1), slab_parity self-check included in the default runs
set -u
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node 8 --master-port 29700 bench.py --gpus 8 --steps 100 --warmup 10 2> $O/r02_bench_n8_chunks.err | grep '^{' > $O/r02_bench_n8_chunks.json
FDTD_B200_P2P_PLANE_SPLIT=1 $TR --nproc-per-node 8 --master-port 29701 bench.py --gpus 8 --steps 100 --warmup 10 --no-slab-parity 2> $O/r02_bench_n8_planesplit.err | grep '^{' > $O/r02_bench_n8_planesplit.json
$TR --nproc-per-node 4 --master-port 29702 bench.py --gpus 4 --steps 100 --warmup 10 2> $O/r02_bench_n4_chunks.err | grep '^{' > $O/r02_bench_n4_chunks.json
FDTD_B200_P2P_PLANE_SPLIT=1 $TR --nproc-per-node 4 --master-port 29703 bench.py --gpus 4 --steps 100 --warmup 10 --no-slab-parity 2> $O/r02_bench_n4_planesplit.err | grep '^{' > $O/r02_bench_n4_planesplit.json
python bench.py --steps 100 --warmup 10 --no-cpu-baseline 2> $O/r02_bench_n1_same_box.err | grep '^{' > $O/r02_bench_n1_same_box.json
for f in n8_chunks n8_planesplit n4_chunks n4_planesplit n1_same_box; do python - <<PY
import json
try:
    d=json.load(open("$O/r02_bench_$f.json"))
    print("$f", round(d["value"],2), "Gcells/s", round(d["ms_per_step"],4), "ms/step frac", round(d["roofline"]["frac"],3), d.get("slab_parity",{}).get("bit_identical"), d["clocks"])
except Exception as e:
    print("$f FAILED", e); print(open("$O/r02_bench_$f.err").read()[-1200:])
PY
done
