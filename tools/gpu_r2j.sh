#!/bin/bash
# Round-2 GPU call J (1 GPU): launch lists of the other networks' steps (where does MobileNetV2's time go?).
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
for m in mobilenetv2 vgg16; do
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2j_launches_$m.csv \
    python bench.py --model $m --steps 2 --warmup 3 --no-cpu > gpurun_out/r2j_ncu_$m.log 2>&1
  echo "ncu $m rc=$?"
done
echo ALLDONE
