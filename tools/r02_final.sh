#!/bin/bash
# the driver's end-of-round sequence on one GPU: tests, smoke, reference arm, own arm
timeout 600 python -m pytest tests -x -q -m gpu --timeout 120 2>&1 | tail -2
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1 | cut -c1-100
timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_final_reference.json 2> gpurun_out/r02_final.err
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_final_own.json 2>> gpurun_out/r02_final.err
python - <<'PY'
import json
for f in ("reference","own"):
    try:
        d=json.load(open(f"gpurun_out/r02_final_{f}.json"))
        print(f, d.get("impl"), round(d["value"],3), d["unit"], "e2e", round(d["e2e"]["value"],3), "frac", d.get("roofline",{}).get("frac"), "cpu_baseline", (d.get("cpu_baseline") or {}).get("value"), "launches", d.get("gpu_launches"), "clocks", d.get("clocks"))
    except Exception as e: print(f, "FAILED", e)
PY
tail -2 gpurun_out/r02_final.err
