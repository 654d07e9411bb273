#!/bin/bash
# After the coalesced squeeze-and-excitation MLP kernels: SE tests, rates.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_models_gpu.py -x -q -m gpu -p no:cacheprovider \
  -k "squeeze or senet or regnety" > gpurun_out/r2y_tests.log 2>&1
echo "tests rc=$?"; tail -30 gpurun_out/r2y_tests.log | cut -c1-600
timeout 600 python tools/model_rates.py regnety_400mf senet18 2>&1 | tail -3
cp gpurun_out/model_rates.json gpurun_out/r2y_model_rates.json 2>/dev/null
echo ALLDONE
