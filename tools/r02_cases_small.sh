#!/bin/bash
set -u
O=gpurun_out; T=${1:-x}
python bench.py --config app --steps 2000 --warmup 20 > $O/r02_case_app_$T.json 2> $O/r02_case_app_$T.err
python bench.py --config 1 --steps 1000 --warmup 20 > $O/r02_case_1_$T.json 2> $O/r02_case_1_$T.err
for c in app 1; do python - <<PY
import json
try:
    d=json.load(open("$O/r02_case_${c}_$T.json"))
    print("$c", "value %.3f e2e %.3f (lag %.3f) us/step %.3f cpu %.3f (1T %.3f) path %s" % (d["value"], d["e2e"]["value"], d["e2e"]["read_lagging_one_step"]["value"], 1e3*d["ms_per_step"], d["cpu_baseline"]["value"], d["cpu_baseline"]["single_thread"]["value"], d["impl_detail"]["step_path"][:110]))
except Exception as e:
    print("$c", "FAILED", e); print(open("$O/r02_case_${c}_$T.err").read()[-1500:])
PY
done
