#!/bin/bash
# round-2 GPU visit: every GPU test file in its own process (a trapped kernel must not take the other files down),
# smoke, bench (ours + reference arm)
TAG=${1:-r2a}; shift
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/${TAG}_smi.txt
: > gpurun_out/${TAG}_pytest.log
for f in tests/test_gpu_tma.py tests/test_gpu_ops.py tests/test_gpu_tma_model.py tests/test_gpu_model.py tests/test_gpu_main.py tests/test_gpu_data.py tests/test_gpu_images.py tests/test_gpu_dist.py; do
  echo "=== $f" >> gpurun_out/${TAG}_pytest.log
  timeout 600 python -m pytest $f -m gpu -q -p no:cacheprovider --timeout 300 -s >> gpurun_out/${TAG}_pytest.log 2>&1
  echo "pytest $f exit $?"
done
grep -E "^(FAILED|ERROR)|passed|failed|=== " gpurun_out/${TAG}_pytest.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -n 3 gpurun_out/${TAG}_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 "$@" > gpurun_out/${TAG}_bench.log 2>&1
echo "bench exit $?"; tail -n 2 gpurun_out/${TAG}_bench.log | cut -c1-1800
timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/${TAG}_bench_ref.log 2>&1
echo "ref exit $?"; tail -n 1 gpurun_out/${TAG}_bench_ref.log | cut -c1-600
