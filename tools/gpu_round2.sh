#!/bin/bash
# tensor-core VQ first (short leash), then the whole suite, path debug, bench
TAG=${1:-r1b}; shift
mkdir -p gpurun_out
timeout 180 python -m pytest tests/test_gpu_ops.py -q -p no:cacheprovider -k "vq" --timeout 120 > gpurun_out/${TAG}_vq.log 2>&1
echo "vq pytest exit $?"; tail -n 15 gpurun_out/${TAG}_vq.log | cut -c1-200
timeout 900 python -m pytest tests -m gpu -q -p no:cacheprovider --timeout 300 > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/${TAG}_pytest.log | cut -c1-200
timeout 300 python tools/debug_paths.py > gpurun_out/${TAG}_paths.log 2>&1; echo "paths exit $?"
timeout 300 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -n 2 gpurun_out/${TAG}_smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 "$@" > gpurun_out/${TAG}_bench.log 2>&1
echo "bench exit $?"; tail -n 2 gpurun_out/${TAG}_bench.log | cut -c1-600
