#!/bin/bash
# ResNeXt-29 (grouped convs, 1x1 stem): the new kernel tests, the whole-model parity tests, the drop-in test, rates.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_models_gpu.py tests/test_dropin_gpu.py -x -q -m gpu -p no:cacheprovider \
  -k "grouped or 1x1_stem or resnext" > gpurun_out/r2s_tests.log 2>&1
echo "tests rc=$?"; tail -25 gpurun_out/r2s_tests.log | cut -c1-400
timeout 600 python tools/model_rates.py resnext29_2x64d resnet50 2>&1 | tail -3
cp gpurun_out/model_rates.json gpurun_out/r2s_model_rates.json 2>/dev/null
echo ALLDONE
