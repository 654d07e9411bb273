#!/bin/bash
# ResNeXt with the grouped convs run group by group on channel slices: parity, drop-in, rates (A/B against the dense form).
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_models_gpu.py tests/test_dropin_gpu.py -x -q -m gpu -p no:cacheprovider \
  -k "resnext or regnetx_200mf or grouped" > gpurun_out/r3g_tests.log 2>&1
echo "tests rc=$?"; tail -30 gpurun_out/r3g_tests.log | cut -c1-700
for gs in 1 0; do
  echo "B2C_GROUP_SLICES=$gs"
  B2C_GROUP_SLICES=$gs timeout 600 python tools/model_rates.py resnext29_2x64d 2>&1 | tail -1
done
cp gpurun_out/model_rates.json gpurun_out/r3g_model_rates.json 2>/dev/null
echo ALLDONE
