#!/bin/bash
# end-of-round scaling evidence on one 8-GPU box with the driver's own configuration (--steps 20 --warmup 5): N = 8, 4, 2, 1 back to back
set -u
O=gpurun_out; mkdir -p $O
python -m pytest tests/test_fused_gpu.py tests/test_narrow_tiles_gpu.py -m gpu -x -q 2>&1 | tail -2
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
python bench.py --steps 20 --warmup 5 2> $O/scale.err | grep '^{' > $O/r02_scale_n1.json
for n in 2 4 8; do
  $TR --nproc-per-node $n --master-port $((29700 + n)) bench.py --gpus $n --steps 20 --warmup 5 2>> $O/scale.err | grep '^{' > $O/r02_scale_n$n.json
done
$TR --nproc-per-node 8 --master-port 29720 bench.py --gpus 8 --steps 20 --warmup 5 --pml 16 --no-slab-parity 2>> $O/scale.err | grep '^{' > $O/r02_scale_n8_pml16.json
$TR --nproc-per-node 8 --master-port 29721 bench.py --gpus 8 --steps 20 --warmup 5 --impl reference 2>> $O/scale.err | grep '^{' > $O/r02_scale_n8_reference.json
for f in n1 n2 n4 n8 n8_pml16 n8_reference; do python - <<PY
import json
try:
    d=json.load(open("$O/r02_scale_$f.json"))
    print("$f", d["config"].get("lattice"), round(d["value"],3), d["unit"], "e2e", round(d["e2e"]["value"],3), round(d["ms_per_step"],4), "ms/step frac", round(d.get("roofline",{}).get("frac",0),3), d.get("slab_parity",{}).get("bit_identical"), d.get("clocks",{}).get("sm_mhz"), d.get("clocks",{}).get("reasons"))
except Exception as e:
    print("$f FAILED", e)
PY
done
tail -3 $O/scale.err
