#!/bin/bash
# What the driver runs at round end, on the current build: the GPU suite, smoke, the reference arm and the default bench.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1800 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider > gpurun_out/final_tests.log 2>&1
echo "tests rc=$?"; tail -6 gpurun_out/final_tests.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 2>/dev/null | cut -c1-400
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err
echo "bench rc=$?"; cut -c1-700 gpurun_out/final_bench.json
echo ALLDONE
