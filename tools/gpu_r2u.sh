#!/bin/bash
# DLA (models/dla.py): four-source Root kernel test, whole-model parity, drop-in test, rate; SimpleDLA regression.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_models_gpu.py tests/test_dropin_gpu.py -x -q -m gpu -p no:cacheprovider \
  -k "concat or dla or round2_kernels" > gpurun_out/r2u_tests.log 2>&1
echo "tests rc=$?"; tail -25 gpurun_out/r2u_tests.log | cut -c1-600
timeout 600 python tools/model_rates.py dla simpledla 2>&1 | tail -3
cp gpurun_out/model_rates.json gpurun_out/r2u_model_rates.json 2>/dev/null
echo ALLDONE
