#!/bin/bash
# round-2 visit: compute-sanitizer over the tensor-core / TMA / grid-barrier kernels at small shapes (memcheck,
# racecheck, synccheck), then a source-level ncu capture of the VQ search at the step's size
TAG=${1:-r2s}
mkdir -p gpurun_out
SEL='test_conv_tma_forward_dgrad_wgrad and (case0 or case4 or case7) or test_conv_bn_tma_train_eval_backward and (case0 or case1) or test_vq_tensor_core_lowest_index_ties and 512 or test_mel_collate_kernel_bit_exact or test_c1_tensor_core_kernels and case0 or test_stft_istft_vs_oracle and 5000 or test_griffin_lim_one_iteration'
for tool in memcheck synccheck racecheck; do
  log=gpurun_out/${TAG}_sanitize_${tool}.log
  timeout 700 compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 9 \
    python -m pytest tests/test_gpu_tma.py tests/test_gpu_ops.py tests/test_gpu_data.py tests/test_gpu_audio.py -m gpu -q -p no:cacheprovider -k "$SEL" > $log 2>&1
  echo "$tool exit $? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY|passed|failed' $log | tr '\n' ' ' | cut -c1-300)"
  grep -E "Invalid|Race reported|hazard|Barrier error|divergent" $log | sort | uniq -c | sort -rn | head -8 | cut -c1-220
done
VQ_N=40960 timeout 300 ncu --set full --import-source on --clock-control none -k regex:vq_search -s 2 -c 1 -f -o /tmp/${TAG}_vq \
   python tools/vq_one.py > gpurun_out/${TAG}_vqone.log 2>&1
echo "ncu vq exit $?"
ncu -i /tmp/${TAG}_vq.ncu-rep --page source --csv > gpurun_out/${TAG}_vq_source.csv 2>/dev/null
ncu -i /tmp/${TAG}_vq.ncu-rep --page details > gpurun_out/${TAG}_vq_details.txt 2>/dev/null
du -sh gpurun_out
