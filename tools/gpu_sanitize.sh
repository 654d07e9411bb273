#!/bin/bash
# compute-sanitizer (memcheck, then racecheck on a smaller set) over the kernels added or changed in round 2, small shapes.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
SEL='(repack or stride2_dgrad_on_cta_pairs or over_concat or fused_finalize or statistics_accumulated or dgrad_with_shortcut or test_conv_dgrad or previous_batchnorm) and not 1024 and not 131072'
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 7 --print-limit 20 \
  python -m pytest tests/test_kernels_gpu.py tests/test_models_gpu.py -m gpu -q -x -p no:cacheprovider -k "$SEL" > gpurun_out/sanitize_memcheck.log 2>&1
echo "memcheck rc=$?"; grep -E "ERROR SUMMARY|passed|failed|Invalid|out of bounds" gpurun_out/sanitize_memcheck.log | tail -8
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 7 --print-limit 10 \
  python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -p no:cacheprovider -k "(repack or fused_finalize or (test_conv_dgrad and True-shape0)) and not 131072" > gpurun_out/sanitize_racecheck.log 2>&1
echo "racecheck rc=$?"; grep -E "RACECHECK SUMMARY|passed|failed|hazard" gpurun_out/sanitize_racecheck.log | tail -6
echo ALLDONE
