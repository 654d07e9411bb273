#!/bin/bash
# Round-2 GPU call E (1 GPU): the pieces changed after call D — repack kernel, main_dist.py --amp — and a bench A/B.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests -m gpu -q -p no:cacheprovider -k "repack or amp_flag or sgd or dropin or network" > gpurun_out/r2e_tests.log 2>&1
echo "tests rc=$?"; tail -12 gpurun_out/r2e_tests.log | cut -c1-400
timeout 600 python bench.py --steps 200 --warmup 20 --no-cpu > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err
echo "bench rc=$?"; python -c "import json; d=json.load(open('gpurun_out/r2e_bench.json')); print(d['value'], d['ms_per_step'], d['clocks'], d['e2e']['value'])"
timeout 600 python bench.py --batch 128 --steps 200 --warmup 20 --no-cpu > gpurun_out/r2e_bench_b128.json 2>/dev/null; python -c "import json; d=json.load(open('gpurun_out/r2e_bench_b128.json')); print('b128', d['value'], d['ms_per_step'])"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'repack|sgd' -c 12 --csv --log-file gpurun_out/r2e_repack.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu > /dev/null 2>&1; grep -E "repack|sgd" gpurun_out/r2e_repack.csv | awk -F'","' '{print $5, $NF}' | cut -c1-120 | tail -6
echo ALLDONE
