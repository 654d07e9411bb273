#!/bin/bash
# Round-2 GPU call C (N GPUs of one box): the multi-GPU legs — bench.py under torch.distributed.run (weak + strong +
# grad_mean_ok), accelerate() under DistributedDataParallel on N ranks — plus the tests changed since call B.
N=${1:-2}
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -m gpu -q -p no:cacheprovider -k "accuracy_curve or simpledla or trainer" > gpurun_out/r2c_tests.log 2>&1
echo "tests rc=$?"; tail -8 gpurun_out/r2c_tests.log | cut -c1-300
cat gpurun_out/top1_curve_hard_task.json | head -60
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
  tools/ddp_accelerate_check.py > gpurun_out/r2c_ddp_check_n$N.log 2>&1
echo "ddp check rc=$?"; tail -4 gpurun_out/r2c_ddp_check_n$N.log | cut -c1-1200
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 \
  bench.py --gpus $N --steps 100 --warmup 10 > gpurun_out/r2c_bench_n$N.json 2> gpurun_out/r2c_bench_n$N.err
echo "bench N=$N rc=$?"; cat gpurun_out/r2c_bench_n$N.json; tail -5 gpurun_out/r2c_bench_n$N.err
timeout 600 python bench.py --steps 100 --warmup 10 --no-cpu > gpurun_out/r2c_bench_n1.json 2> gpurun_out/r2c_bench_n1.err
echo "bench N=1 rc=$?"; cat gpurun_out/r2c_bench_n1.json
echo ALLDONE
