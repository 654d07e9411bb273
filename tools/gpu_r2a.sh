#!/bin/bash
# Round-2 GPU call A: full GPU test suite, bench lines (all four models), drop-in rates, ncu launch list and
# --set full capture of one step's conv kernels, tensor-pipe metric survey on the MMA probe.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r2a_nvsmi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2a_tests.log 2>&1
echo "tests rc=$?" | tee -a gpurun_out/r2a_tests.log
tail -40 gpurun_out/r2a_tests.log
B2C_FUSED_FINALIZE=1 timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -x --deselect tests/test_reference_scripts_gpu.py > gpurun_out/r2a_tests_fused.log 2>&1
echo "tests(fused finalize) rc=$?" | tee -a gpurun_out/r2a_tests_fused.log
tail -15 gpurun_out/r2a_tests_fused.log
for f in 0 1; do for b in 1024 128; do
  B2C_FUSED_FINALIZE=$f timeout 600 python bench.py --steps 100 --warmup 10 --no-cpu --batch $b > gpurun_out/r2a_ab_f${f}_b${b}.json 2> gpurun_out/r2a_ab_f${f}_b${b}.err
  echo "A/B fused=$f batch=$b rc=$?"; python -c "import json,sys; d=json.load(open('gpurun_out/r2a_ab_f${f}_b${b}.json')); print(d['value'], d['ms_per_step'], d['launches_per_step'], d['clocks'], d['final_loss'])"
done; done
timeout 600 python bench.py --steps 200 --warmup 20 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err
echo "bench rc=$?"; cat gpurun_out/r2a_bench.json; tail -5 gpurun_out/r2a_bench.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2a_ref.json 2> gpurun_out/r2a_ref.err
cat gpurun_out/r2a_ref.json
for m in vgg16 preact18 mobilenetv2; do
  timeout 600 python bench.py --model $m --steps 50 --warmup 5 --no-cpu > gpurun_out/r2a_bench_$m.json 2> gpurun_out/r2a_bench_$m.err
  echo "bench $m rc=$?"; cat gpurun_out/r2a_bench_$m.json; tail -3 gpurun_out/r2a_bench_$m.err
done
timeout 600 python tools/dropin_rate.py > gpurun_out/r2a_dropin.log 2>&1
echo "dropin rc=$?"; tail -4 gpurun_out/r2a_dropin.log
# --- ncu: launch list of the bench command (2 steps), then --set full of one eager step's conv kernels
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r2a_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2a_ncu_launch.log 2>&1
echo "ncu launches rc=$?"
timeout 1500 ncu --profile-from-start off --set full --clock-control none --import-source on \
  -k regex:'conv_halo|tap_gemm|wgrad' -f -o gpurun_out/r2a_conv_step python tools/ncu_step.py > gpurun_out/r2a_ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -3 gpurun_out/r2a_ncu_full.log
ls -la gpurun_out/r2a_conv_step.ncu-rep
# --- which counter tracks tcgen05.mma? every metric with tensor / umma in its name, read on the MMA probe
ncu --query-metrics 2>/dev/null | grep -iE 'tensor|umma|utc' > gpurun_out/r2a_tensor_metric_names.txt
wc -l gpurun_out/r2a_tensor_metric_names.txt
timeout 600 ncu --clock-control none --metrics "regex:.*tensor.*,sm__cycles_elapsed.max,sm__cycles_active.max" --csv \
  --log-file gpurun_out/r2a_probe_tensor_metrics.csv tools/bin/umma2_probe > gpurun_out/r2a_probe.log 2>&1
echo "probe ncu rc=$?"; tail -12 gpurun_out/r2a_probe.log
echo ALLDONE
