#!/bin/bash
# Round-2 GPU call K (1 GPU): BatchNorm apply kernels on uniform grids for every channel count — tests, MobileNetV2 /
# ResNet-50 rates before/after knobs, ResNet-18 sanity.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -k "bn_ or models or dropin or network" > gpurun_out/r2k_tests.log 2>&1
echo "tests rc=$?"; tail -8 gpurun_out/r2k_tests.log | cut -c1-300
for f in 0 1; do
  B2C_FUSED_FINALIZE=$f timeout 600 python bench.py --model mobilenetv2 --steps 100 --warmup 10 --no-cpu > gpurun_out/r2k_mbv2_f$f.json 2>/dev/null
  python -c "import json; d=json.load(open('gpurun_out/r2k_mbv2_f$f.json')); print('mobilenetv2 fused_finalize=$f', d['value'], d['ms_per_step'], d['launches_per_step'], d['roofline']['frac'], d['clocks']['sm_mhz'])"
done
timeout 600 python tools/model_rates.py resnet50 simpledla vgg16 preact_resnet18 2>&1 | grep '^{'
timeout 600 python bench.py --steps 100 --warmup 10 --no-cpu | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][0]); print('resnet18', d['value'], d['ms_per_step'], d['clocks']['sm_mhz'])"
echo ALLDONE
