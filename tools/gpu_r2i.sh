#!/bin/bash
# Round-2 GPU call I (8 GPUs): gradient bucket size on the strong-scaling configuration (128 images per GPU).
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
for mb in 25 8 4 2 1; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29581 \
    bench.py --gpus 8 --batch 128 --no-strong --bucket-mb $mb --steps 200 --warmup 20 > gpurun_out/r2i_b128_mb$mb.out 2> gpurun_out/r2i_b128_mb$mb.err
  echo "bucket $mb MiB rc=$?"; grep '^{' gpurun_out/r2i_b128_mb$mb.out | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('   ', round(d['value']), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value']), d['clocks']['sm_mhz'], d.get('grad_mean_ok'))"
done
for mb in 25 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29582 \
    bench.py --gpus 8 --no-strong --bucket-mb $mb --steps 100 --warmup 10 > gpurun_out/r2i_b1024_mb$mb.out 2> gpurun_out/r2i_b1024_mb$mb.err
  echo "weak bucket $mb MiB rc=$?"; grep '^{' gpurun_out/r2i_b1024_mb$mb.out | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('   ', round(d['value']), round(d['ms_per_step'],4))"
done
echo ALLDONE
