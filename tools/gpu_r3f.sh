#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -p no:cacheprovider -k "channel_slices" > gpurun_out/r3f_tests.log 2>&1
echo "tests rc=$?"; tail -30 gpurun_out/r3f_tests.log | cut -c1-700
echo ALLDONE
