#!/bin/bash
# programmatic dependent launch of the step kernels: A/B against plain launches (FDTD_B200_NO_PDL=1), same box, alternating
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for spec in "3 256,256,256" "3 512,512,512" "thin 1000,1000,1" "thin 2048,2048,1" "thin 4096,4096,1" "3 320,320,320"; do
  set -- $spec
  for rep in 1 2; do for nopdl in 1 0; do
  FDTD_B200_NO_PDL=$nopdl python bench.py --config $1 --lattice $2 --steps 50 --warmup 5 --passes 9 --no-cpu-baseline --no-slab-parity 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$2 no_pdl=$nopdl value %.2f e2e %.2f frac %.3f us/step %.2f' % (d['value'], d['e2e']['value'], d['roofline']['frac'], 1e3*d['ms_per_step']))"
  done; done
done
