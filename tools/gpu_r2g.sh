#!/bin/bash
# Round-2 GPU call G (2 GPUs): how many SMs NCCL may take while the persistent conv kernels run — NCCL_MAX_CTAS A/B on
# the weak-scaling step (1024 images per GPU) and the strong-scaling leg (512 per GPU).
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
for cta in default 2 4 8 16; do
  if [ $cta = default ]; then unset NCCL_MAX_CTAS; else export NCCL_MAX_CTAS=$cta; fi
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 \
    bench.py --gpus 2 --steps 100 --warmup 10 > gpurun_out/r2g_bench_cta_$cta.out 2> gpurun_out/r2g_bench_cta_$cta.err
  echo "NCCL_MAX_CTAS=$cta rc=$?"; grep '^{' gpurun_out/r2g_bench_cta_$cta.out | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); s=d.get('strong') or {}
print('   weak', round(d['value']), round(d['ms_per_step'],4), 'strong', round(s.get('value',0)), s.get('ms_per_step'), 'clk', d['clocks']['sm_mhz'])"
done
unset NCCL_MAX_CTAS
timeout 600 python bench.py --gpus 1 --steps 100 --warmup 10 --no-cpu | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][0]); print('N=1', round(d['value']), round(d['ms_per_step'],4), d['clocks']['sm_mhz'])"
echo ALLDONE
