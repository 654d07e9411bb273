#!/bin/bash
# Multi-GPU regression on the final build (N GPUs of one box): accelerate() under DDP, bench weak + strong, reference arm.
N=${1:-2}
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29543 \
  tools/ddp_accelerate_check.py > gpurun_out/r3c_ddp_check_n$N.log 2>&1
echo "ddp check rc=$?"; tail -3 gpurun_out/r3c_ddp_check_n$N.log | cut -c1-900
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 \
  bench.py --gpus $N --steps 50 --warmup 10 > gpurun_out/r3c_bench_n$N.json 2> gpurun_out/r3c_bench_n$N.err
echo "bench N=$N rc=$?"; cut -c1-1500 gpurun_out/r3c_bench_n$N.json; tail -3 gpurun_out/r3c_bench_n$N.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29545 \
  bench.py --impl reference --gpus $N --steps 5 --warmup 2 2>/dev/null | cut -c1-300
echo ALLDONE
