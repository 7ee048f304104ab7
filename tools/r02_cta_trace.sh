#!/bin/bash
# per-CTA (SM, start, end) stamps of the fused kernel's last launch, with a diagnostic build of the library (-DFDTD_B200_CTA_TRACE)
for spec in ${SPECS:-"256,256,256 64" "256,256,256 16" "512,512,512 64"}; do
  set -- $spec
  FDTD_B200_TUNE=0,0,$2 FDTD_B200_CTA_TRACE_FILE=gpurun_out/cta_trace_${1//,/x}_zc$2.txt python bench.py --lattice $1 --steps 10 --warmup 3 --passes 5 --no-cpu-baseline --no-slab-parity 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$1 zc $2 value %.2f ms %.4f' % (d['value'], d['ms_per_step']))"
done
ls -la gpurun_out/cta_trace_*
