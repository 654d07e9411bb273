#!/bin/bash
# SENet18: squeeze-and-excitation kernel test, whole-model parity, drop-in test, rate; PreActResNet regression.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_models_gpu.py tests/test_dropin_gpu.py -x -q -m gpu -p no:cacheprovider \
  -k "squeeze or senet or preact" > gpurun_out/r2v_tests.log 2>&1
echo "tests rc=$?"; tail -30 gpurun_out/r2v_tests.log | cut -c1-600
timeout 600 python tools/model_rates.py senet18 preact_resnet18 2>&1 | tail -3
cp gpurun_out/model_rates.json gpurun_out/r2v_model_rates.json 2>/dev/null
echo ALLDONE
