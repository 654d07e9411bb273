#!/bin/bash
# Round-2 GPU call L (1 GPU): epilogue operand requested one tile ahead in conv_halo2 (64-wide tiles) — tests, bench,
# launch list.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -k "conv or network or models or dropin" > gpurun_out/r2l_tests.log 2>&1
echo "tests rc=$?"; tail -8 gpurun_out/r2l_tests.log | cut -c1-300
timeout 600 python bench.py --steps 200 --warmup 20 --no-cpu > gpurun_out/r2l_bench.json 2> gpurun_out/r2l_bench.err
echo "bench rc=$?"; python -c "import json; d=json.load(open('gpurun_out/r2l_bench.json')); r=d['roofline']; print(d['value'], d['ms_per_step'], d['clocks'], d['e2e']['value'], 'halo', r['achieved'], r['ms_per_step'], 'tap', r['tap_gemm']['achieved'], 'wgrad', r['wgrad']['achieved'])"
timeout 600 python bench.py --batch 128 --steps 200 --warmup 20 --no-cpu > gpurun_out/r2l_bench_b128.json 2>/dev/null; python -c "import json; d=json.load(open('gpurun_out/r2l_bench_b128.json')); print('b128', d['value'], d['ms_per_step'])"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r2l_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2l_ncu_launch.log 2>&1
echo "ncu launches rc=$?"; grep "conv_halo2_kernel<64" gpurun_out/r2l_launches.csv | awk -F'","' '{print $5, $NF}' | sed 's/(CUtensorMap.*)//' | tail -8
echo ALLDONE
