#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_models_gpu.py -x -q -m gpu -p no:cacheprovider -k "resnext29_4x64d" > gpurun_out/r3h_tests.log 2>&1
echo "tests rc=$?"; tail -30 gpurun_out/r3h_tests.log | cut -c1-700
for gs in 1 0; do
  echo "B2C_GROUP_SLICES=$gs"
  B2C_GROUP_SLICES=$gs timeout 600 python tools/model_rates.py resnext29_4x64d 2>&1 | tail -1
done
echo ALLDONE
