#!/bin/bash
# RegNetY_400MF (squeeze-and-excitation inside the bottleneck block): parity, rates; bottleneck-family regression.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_models_gpu.py -x -q -m gpu -p no:cacheprovider \
  -k "regnet or resnet50 or resnext29_2x64d" > gpurun_out/r2w_tests.log 2>&1
echo "tests rc=$?"; tail -30 gpurun_out/r2w_tests.log | cut -c1-600
timeout 600 python tools/model_rates.py regnety_400mf regnetx_400mf 2>&1 | tail -3
cp gpurun_out/model_rates.json gpurun_out/r2w_model_rates.json 2>/dev/null
echo ALLDONE
