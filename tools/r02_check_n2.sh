#!/bin/bash
# after the kernel changes of the round: multi-GPU parity tests on 2 GPUs, N=2 bench (plain and with CPML layers), N=1 on the same box
set -u
O=gpurun_out; mkdir -p $O
python -m pytest tests/test_multigpu.py -m gpu -x -q 2>&1 | tail -6 > $O/r02_pytest_n2_b.log; tail -3 $O/r02_pytest_n2_b.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29600"
$TR bench.py --gpus 2 --steps 20 --warmup 5 2> $O/n2.err | grep '^{' > $O/r02_bench_n2_final.json
$TR bench.py --gpus 2 --steps 20 --warmup 5 --pml 16 --no-slab-parity 2>> $O/n2.err | grep '^{' > $O/r02_bench_n2_pml16.json
python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>> $O/n2.err | grep '^{' > $O/r02_bench_n1_same_box_b.json
python bench.py --steps 20 --warmup 5 --pml 16 --lattice 512,512,1024 --no-cpu-baseline 2>> $O/n2.err | grep '^{' > $O/r02_bench_n1_pml16_tall.json
for f in n2_final n2_pml16 n1_same_box_b n1_pml16_tall; do python - <<PY
import json
try:
    d=json.load(open("$O/r02_bench_$f.json"))
    print("$f", d["config"]["lattice"], round(d["value"],2), "Gcells/s e2e", round(d["e2e"]["value"],2), round(d["ms_per_step"],4), "ms/step frac", round(d["roofline"]["frac"],3), d.get("slab_parity",{}).get("bit_identical"), d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
except Exception as e:
    print("$f FAILED", e)
PY
done
tail -5 $O/n2.err
