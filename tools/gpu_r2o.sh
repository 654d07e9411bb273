#!/bin/bash
# Round-2 GPU call O (1 GPU): depthwise conv fed by the previous BatchNorm (MobileNetV2) — tests and A/B.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -k "depthwise or mobilenetv2 or models" > gpurun_out/r2o_tests.log 2>&1
echo "tests rc=$?"; tail -8 gpurun_out/r2o_tests.log | cut -c1-400
B2C_DW_FUSED_INPUT=0 timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -k "mobilenetv2" 2>&1 | tail -2
for f in 0 1; do
  B2C_DW_FUSED_INPUT=$f timeout 600 python bench.py --model mobilenetv2 --steps 100 --warmup 10 --no-cpu > gpurun_out/r2o_mbv2_f$f.json 2>/dev/null
  python -c "import json; d=json.load(open('gpurun_out/r2o_mbv2_f$f.json')); print('mobilenetv2 dw_fused_input=$f', d['value'], d['ms_per_step'], d['launches_per_step'], d['roofline']['frac'], d['clocks']['sm_mhz'], d['final_loss'])"
done
echo ALLDONE
