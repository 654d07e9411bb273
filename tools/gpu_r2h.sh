#!/bin/bash
# Round-2 GPU call H (1 GPU): stride-2 / fused-shortcut data gradients on CTA pairs — tests, bench, launch list.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -k "dgrad or shortcut or network or models or dropin or stride2" > gpurun_out/r2h_tests.log 2>&1
echo "tests rc=$?"; tail -12 gpurun_out/r2h_tests.log | cut -c1-400
timeout 600 python bench.py --steps 200 --warmup 20 --no-cpu > gpurun_out/r2h_bench.json 2> gpurun_out/r2h_bench.err
echo "bench rc=$?"; python -c "import json; d=json.load(open('gpurun_out/r2h_bench.json')); r=d['roofline']; print(d['value'], d['ms_per_step'], d['clocks'], d['e2e']['value'], 'tap', r['tap_gemm']['achieved'], r['tap_gemm']['ms_per_step'], 'halo', r['achieved'], 'wgrad', r['wgrad']['achieved'])"
B2C_CTA_PAIR=halo timeout 600 python bench.py --steps 200 --warmup 20 --no-cpu > gpurun_out/r2h_bench_pairhalo.json 2>/dev/null; python -c "import json; d=json.load(open('gpurun_out/r2h_bench_pairhalo.json')); r=d['roofline']; print('B2C_CTA_PAIR=halo', d['value'], d['ms_per_step'], d['clocks']['sm_mhz'], 'tap', r['tap_gemm']['achieved'], r['tap_gemm']['ms_per_step'])"
timeout 600 python bench.py --batch 128 --steps 200 --warmup 20 --no-cpu > gpurun_out/r2h_bench_b128.json 2>/dev/null; python -c "import json; d=json.load(open('gpurun_out/r2h_bench_b128.json')); print('b128', d['value'], d['ms_per_step'])"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r2h_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2h_ncu_launch.log 2>&1
echo "ncu launches rc=$?"
echo ALLDONE
