#!/bin/bash
# A/B of the second MMA-issuing thread in wgrad_tma_kernel (NSG_WGRAD_ISSUERS)
TAG=${1:-r2v}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ops.py tests/test_gpu_tma.py tests/test_gpu_tma_model.py -x -m gpu -q -p no:cacheprovider > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; tail -n 3 gpurun_out/${TAG}_pytest.log | cut -c1-300; grep -E "^(FAILED|ERROR)|hang report\]|^E  " gpurun_out/${TAG}_pytest.log | head -10 | cut -c1-300
for n in 2 1 2 1; do
  NSG_WGRAD_ISSUERS=$n timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_bench_i$n.log 2>&1
  echo "wgrad issuers=$n bench exit $?"; tail -n 1 gpurun_out/${TAG}_bench_i$n.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['us'], d['roofline']['frac'])" 2>&1 | tail -n 1 | cut -c1-300
done
for n in 2 1; do
  NSG_WGRAD_ISSUERS=$n timeout 200 python tools/bench_ops.py > gpurun_out/${TAG}_ops_i$n.log 2>&1; echo "wgrad issuers=$n ops exit $?"
  grep -E "wgrad" gpurun_out/${TAG}_ops_i$n.log | cut -c1-130 | head -24
  NSG_WGRAD_ISSUERS=$n timeout 120 python tools/timeline_wgrad.py > gpurun_out/${TAG}_timeline_i$n.log 2>&1; echo "timeline exit $?"; tail -n 8 gpurun_out/${TAG}_timeline_i$n.log | cut -c1-300
done
