mkdir -p gpurun_out
: > gpurun_out/each.log
for k in "size0" "size1" "size2" "size3" "size4" "size5" "size6" "size7" "size8" "size9" "size10" "probes_sources" "on_slabs"; do
  timeout 120 python -m pytest tests/test_bulk_gpu.py -m gpu -q -k "$k" > gpurun_out/each_$k.log 2>&1
  echo "$k rc=$? $(tail -1 gpurun_out/each_$k.log)" >> gpurun_out/each.log
done
cat gpurun_out/each.log
