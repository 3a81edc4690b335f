#!/bin/bash
# compute-sanitizer over the tcgen05 / TMA / grid-barrier kernels at small shapes (sanitizers).
# Each tool runs a handful of op-level tests in its own process; summaries land in gpurun_out/<tag>_sanitize_*.log
# (copy the tails into profiles/).  Usage: bash tools/gpu_sanitize.sh <tag>
TAG=${1:-r2s}
mkdir -p gpurun_out
SEL='test_conv_tma_forward_dgrad_wgrad and (case0 or case2 or case4 or case7) or test_conv_bn_tma_train_eval_backward and (case0 or case1) or test_vq_tensor_core_lowest_index_ties and 512 or test_mel_collate_kernel_bit_exact'
for tool in memcheck racecheck synccheck; do
  log=gpurun_out/${TAG}_sanitize_${tool}.log
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 9 \
    python -m pytest tests/test_gpu_tma.py tests/test_gpu_ops.py tests/test_gpu_data.py -m gpu -q -p no:cacheprovider -x -k "$SEL" > $log 2>&1
  echo "$tool exit $? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY|passed|failed' $log | tr '\n' ' ' | cut -c1-300)"
done
