#!/bin/bash
# Bench lines of BASELINE configs 3-5 on the final build.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
for m in vgg16 preact18 mobilenetv2; do
  timeout 600 python bench.py --model $m --steps 50 --warmup 10 > gpurun_out/r3j_bench_$m.json 2> gpurun_out/r3j_bench_$m.err
  echo "$m rc=$?"; grep '^{' gpurun_out/r3j_bench_$m.json | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(d['config'].get('model'), round(d['value']), d['ms_per_step'], 'e2e', round(d['e2e']['value']), d['roofline']['bound'], round(d['roofline']['frac'],3), d['clocks'])"
done
echo ALLDONE
