#!/bin/bash
# the lattices the reference really runs: one bench line each, CPU leg on the same lattice
set -u
O=gpurun_out; T=${1:-base}
mkdir -p $O
python bench.py --config app --steps 2000 --warmup 20 > $O/r02_case_app_$T.json 2> $O/r02_case_app_$T.err
python bench.py --config 1 --steps 1000 --warmup 20 > $O/r02_case_1_$T.json 2> $O/r02_case_1_$T.err
python bench.py --config thin --steps 50 --warmup 5 > $O/r02_case_thin_$T.json 2> $O/r02_case_thin_$T.err
python bench.py --config 2 --steps 100 --warmup 10 > $O/r02_case_2_$T.json 2> $O/r02_case_2_$T.err
for c in app 1 thin 2; do python - <<PY
import json
try:
    d=json.load(open("$O/r02_case_${c}_$T.json"))
    print("$c", "value %.3f e2e %.3f ms/step %.5f cpu %.3f (1T %.3f) frac %.3f path %s" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["cpu_baseline"]["value"], d["cpu_baseline"]["single_thread"]["value"], d["roofline"]["frac"], d["impl_detail"]["step_path"]))
except Exception as e:
    print("$c", "FAILED", e); print(open("$O/r02_case_${c}_$T.err").read()[-1500:])
PY
done
