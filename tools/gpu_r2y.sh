#!/bin/bash
# quick check of a kernel change: op tests, the bench line, VQ timings
TAG=${1:-r2y}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops.py tests/test_gpu_tma.py tests/test_gpu_tma_model.py -x -m gpu -q -p no:cacheprovider > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; tail -n 3 gpurun_out/${TAG}_pytest.log | cut -c1-300; grep -E "^(FAILED|ERROR)|hang report\]|^E  " gpurun_out/${TAG}_pytest.log | head -10 | cut -c1-300
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_bench.log 2>&1
echo "bench exit $?"; tail -n 1 gpurun_out/${TAG}_bench.log | cut -c1-1900
timeout 300 python tools/bench_vq.py > gpurun_out/${TAG}_vq.log 2>&1; echo "vq exit $?"; grep "vq_search" gpurun_out/${TAG}_vq.log | grep "mode=1" | cut -c1-160
timeout 200 python tools/bench_ops.py > gpurun_out/${TAG}_ops.log 2>&1; echo "ops exit $?"; grep -E "c1t|fwd\+stats  warm|dgrad\+mask warm|wgrad      warm" gpurun_out/${TAG}_ops.log | cut -c1-120
