#!/bin/bash
# does an idle gap between timed passes keep a long (K=200) run out of the power cap?  same box, gaps 0 / 100 / 300 / 1000 ms
for g in 0 100 300 1000 0; do
  python bench.py --steps 200 --warmup 10 --no-cpu-baseline --pass-gap-ms $g 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('gap $g ms: value %.2f (best %.2f worst %.2f) e2e %.2f frac %.4f clocks %s' % (d['value'], d['timing']['best_value'], d['timing']['worst_value'], d['e2e']['value'], d['roofline']['frac'], d['clocks']))"
done
