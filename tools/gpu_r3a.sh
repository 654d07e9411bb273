#!/bin/bash
# GoogLeNet: whole-model parity, drop-in test, rate.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_models_gpu.py tests/test_dropin_gpu.py -x -q -m gpu -p no:cacheprovider \
  -k "googlenet" > gpurun_out/r3a_tests.log 2>&1
echo "tests rc=$?"; tail -40 gpurun_out/r3a_tests.log | cut -c1-700
timeout 600 python tools/model_rates.py googlenet 2>&1 | tail -3
cp gpurun_out/model_rates.json gpurun_out/r3a_model_rates.json 2>/dev/null
echo ALLDONE
