#!/bin/bash
# Round-2 GPU call B: new tests (SimpleDLA, fused-finalize entry points), the suite under B2C_FUSED_FINALIZE=1, hard-task
# calibration, ncu launch lists (batch 1024 and 128) and the --set full capture summarised ON the box (the raw report
# exceeds what gpurun copies back; a trimmed report of the first launches is kept for the source view).
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -k "dla or DLA or concat or accumulated or fused_finalize or depthwise or references_own or main_py" > gpurun_out/r2b_tests_new.log 2>&1
echo "new tests rc=$?"; tail -30 gpurun_out/r2b_tests_new.log
B2C_FUSED_FINALIZE=1 timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider --deselect tests/test_reference_scripts_gpu.py > gpurun_out/r2b_tests_fused.log 2>&1
echo "tests(fused finalize) rc=$?"; tail -40 gpurun_out/r2b_tests_fused.log | cut -c1-300
timeout 900 python tools/calibrate_hard_task.py 2 3 4.5 6 > gpurun_out/r2b_calib.log 2>&1; tail -5 gpurun_out/r2b_calib.log
for b in 1024 128; do
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r2b_launches_b$b.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --batch $b > gpurun_out/r2b_ncu_launch_b$b.log 2>&1
  echo "ncu launches b$b rc=$?"
done
timeout 1500 ncu --profile-from-start off --set full --clock-control none \
  -k regex:'conv_halo|tap_gemm|wgrad' -f -o /tmp/r2b_conv_step python tools/ncu_step.py > gpurun_out/r2b_ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/r2b_ncu_full.log
python tools/ncu_summary.py /tmp/r2b_conv_step.ncu-rep gpurun_out/r2b_ncu_conv_kernels.json 1024 | tee gpurun_out/r2b_ncu_summary.txt
ncu -i /tmp/r2b_conv_step.ncu-rep --page raw --csv > /tmp/raw.csv 2>/dev/null; wc -c /tmp/raw.csv
python - <<'PY'
import csv
rows = list(csv.reader(open('/tmp/raw.csv')))
hdr = next(i for i, r in enumerate(rows) if 'Kernel Name' in r)
keep = [i for i, n in enumerate(rows[hdr]) if n in ('ID', 'Kernel Name', 'Grid Size', 'Block Size') or any(k in n for k in (
    'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'xbar2l1tex_read_bytes.sum', 'tensor', 'umma',
    'sm__throughput.avg.pct', 'dram__throughput.avg.pct', 'lts__t_bytes.sum', 'sm__cycles_elapsed.max', 'sm__cycles_active.avg',
    'smsp__inst_executed.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'sm__inst_executed_pipe'))]
with open('gpurun_out/r2b_ncu_conv_step_raw_selected.csv', 'w', newline='') as f:
    w = csv.writer(f)
    for r in rows[hdr:]:
        if len(r) == len(rows[hdr]):
            w.writerow([r[i] for i in keep])
PY
ls -la gpurun_out/r2b_ncu_conv_step_raw_selected.csv
# a small report with source correlation for the three dominant kernels (first launch of each)
for k in conv_halo2_kernel tap_gemm2_kernel wgrad_row2_kernel; do
  timeout 600 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:$k -c 1 -f \
    -o gpurun_out/r2b_src_$k python tools/ncu_step.py > /dev/null 2>&1; echo "src $k rc=$?"
done
ncu --query-metrics 2>/dev/null | grep -iE 'tensor|umma|utc' > gpurun_out/r2b_tensor_metric_names.txt; wc -l gpurun_out/r2b_tensor_metric_names.txt
timeout 600 ncu --clock-control none --metrics "regex:.*tensor.*,sm__cycles_elapsed.max,sm__cycles_active.max" --csv \
  --log-file gpurun_out/r2b_probe_tensor_metrics.csv tools/bin/umma2_probe > gpurun_out/r2b_probe.log 2>&1
echo "probe ncu rc=$?"
du -sh gpurun_out
echo ALLDONE
