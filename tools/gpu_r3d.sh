#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests/test_reference_scripts_gpu.py -x -q -m gpu -p no:cacheprovider -k "main_py" > gpurun_out/r3d_tests.log 2>&1
echo "tests rc=$?"; tail -30 gpurun_out/r3d_tests.log | cut -c1-900
echo ALLDONE
