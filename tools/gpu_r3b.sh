#!/bin/bash
# PreActResNet-50 (PreActBottleneck): parity, rate; PreActResNet-18 regression.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_models_gpu.py -x -q -m gpu -p no:cacheprovider -k "preact" > gpurun_out/r3b_tests.log 2>&1
echo "tests rc=$?"; tail -30 gpurun_out/r3b_tests.log | cut -c1-700
timeout 600 python tools/model_rates.py preact_resnet50 resnet50 2>&1 | tail -3
cp gpurun_out/model_rates.json gpurun_out/r3b_model_rates.json 2>/dev/null
echo ALLDONE
