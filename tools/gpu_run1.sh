#!/bin/bash
# First GPU bring-up: hardware probe, then each kernel test group in its own process (a trapped
# kernel poisons its CUDA context; the next group still runs).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/nvsmi.txt 2>&1
timeout 120 ./tools/bin/umma_probe > gpurun_out/probe.log 2>&1
echo "probe exit $?" >> gpurun_out/probe.log
tail -45 gpurun_out/probe.log
for t in test_conv_fprop test_conv_fprop_block_n test_conv_fprop_multi_wave test_conv_dgrad test_conv_wgrad \
         test_conv_wgrad_large_k_split test_stem_im2col_and_conv test_bn_forward_chain test_bn_relu_backward \
         test_head_and_loss test_sgd_step test_repack_weights; do
  echo "=== $t" | tee -a gpurun_out/tests.log
  timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "$t and not ${t}_" -p no:cacheprovider 2>&1 | tail -25 | tee -a gpurun_out/tests.log
done
echo ALLDONE
