#!/bin/bash
# Final-build scaling run (N GPUs of one box): bench.py at N = 1, 2, 4, 8 back to back (weak scaling `value` + the `strong`
# leg of each N > 1 line), as the driver runs it, and accelerate() under DDP on N ranks.
NMAX=${1:-8}
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
for N in 1 2 4 8; do
  [ $N -gt $NMAX ] && continue
  if [ $N -eq 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 100 --warmup 10 --no-cpu > gpurun_out/r3e_bench_n1.out 2> gpurun_out/r3e_bench_n1.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29540+N)) \
      bench.py --gpus $N --steps 100 --warmup 10 > gpurun_out/r3e_bench_n$N.out 2> gpurun_out/r3e_bench_n$N.err
  fi
  echo "bench N=$N rc=$?"; grep '^{' gpurun_out/r3e_bench_n$N.out | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
s=d.get('strong') or {}
print('N', d['n_gpus'], 'weak', round(d['value']), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), 'grad_mean_ok', d.get('grad_mean_ok'), 'in_sync', d.get('replicas_in_sync'), '| strong', s.get('value'), s.get('ms_per_step'), s.get('per_gpu_batch'), (s.get('grad_mean_ok') or {}).get('ok'), 'clocks', d['clocks'])"
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NMAX --master-addr 127.0.0.1 --master-port 29561 \
  tools/ddp_accelerate_check.py > gpurun_out/r3e_ddp_check_n$NMAX.log 2>&1
echo "ddp check rc=$?"; grep '^{' gpurun_out/r3e_ddp_check_n$NMAX.log | cut -c1-900
echo ALLDONE
