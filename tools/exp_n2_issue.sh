# N=2 (peer-memory halos): first plane requests at kernel start (default) vs after the scans (FDTD_B200_LATE_ISSUE=1), alternating.
for rep in 1 2; do
for A in 0 1; do
  FDTD_B200_LATE_ISSUE=$A python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2959$A bench.py --gpus 2 --steps 200 --warmup 5 --no-cpu-baseline 2>> gpurun_out/n2_issue.err | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N=2 late_issue=$A value', round(d['value'],2), 'ms', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value'],2), 'lagging', round(d['e2e']['read_lagging_one_step']['value'],2))" >> gpurun_out/exp_n2_issue.txt
done
done
cat gpurun_out/exp_n2_issue.txt
