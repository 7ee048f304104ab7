#!/bin/bash
# where the resident kernel spends a step: clock64 stamps of thread 0 of every CTA (FDTD_B200_RESIDENT_DEBUG=1)
for cfg in app 1; do
  for mode in 0 1; do
    echo "== config $cfg barrier_mode $mode"
    FDTD_B200_RESIDENT_DEBUG=1 FDTD_B200_RESIDENT_BARRIER=$mode python bench.py --config $cfg --steps 50 --warmup 5 --passes 1 --no-cpu-baseline 2>&1 >/dev/null | grep "resident debug" | sort | uniq -c | sort -rn | head -3
  done
done
