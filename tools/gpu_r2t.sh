#!/bin/bash
# RegNetX (grouped convs over padded channel counts): whole-model parity, drop-in test, rates.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_models_gpu.py tests/test_dropin_gpu.py -x -q -m gpu -p no:cacheprovider \
  -k "regnet or resnext or grouped or padded_channel_layout" > gpurun_out/r2t_tests.log 2>&1
echo "tests rc=$?"; tail -25 gpurun_out/r2t_tests.log | cut -c1-600
timeout 600 python tools/model_rates.py regnetx_200mf regnetx_400mf resnext29_2x64d 2>&1 | tail -4
cp gpurun_out/model_rates.json gpurun_out/r2t_model_rates.json 2>/dev/null
echo ALLDONE
