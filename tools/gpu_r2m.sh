#!/bin/bash
# Round-2 GPU call M (final build): whole GPU suite on the current build, smoke, the default bench line, model rates (ResNet-50, SimpleDLA
# included), ncu launch list of the bench command.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1800 python -m pytest tests -m gpu -q -p no:cacheprovider -x > gpurun_out/r2m_tests.log 2>&1
echo "tests rc=$?"; tail -12 gpurun_out/r2m_tests.log | cut -c1-300
timeout 300 python __graft_entry__.py smoke > gpurun_out/r2m_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2m_smoke.log
timeout 600 python bench.py > gpurun_out/r2m_bench.json 2> gpurun_out/r2m_bench.err
echo "bench rc=$?"; cat gpurun_out/r2m_bench.json; tail -3 gpurun_out/r2m_bench.err
timeout 600 python bench.py --batch 128 --steps 100 --warmup 10 --no-cpu > gpurun_out/r2m_bench_b128.json 2>/dev/null; python -c "import json; d=json.load(open('gpurun_out/r2m_bench_b128.json')); print('b128', d['value'], d['ms_per_step'])"
timeout 600 python tools/model_rates.py resnet50 simpledla > gpurun_out/r2m_model_rates.log 2>&1; tail -3 gpurun_out/r2m_model_rates.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r2m_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2m_ncu_launch.log 2>&1
echo "ncu launches rc=$?"
echo ALLDONE
