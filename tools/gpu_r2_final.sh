#!/bin/bash
# round-end check exactly as the driver runs it: the GPU suite in ONE process with -x, smoke(), bench (both arms)
TAG=${1:-r2z}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest (one process, -x) exit $?"; tail -n 4 gpurun_out/${TAG}_pytest.log | cut -c1-300
grep -E "^(FAILED|ERROR)|hang report\]|^E  " gpurun_out/${TAG}_pytest.log | cut -c1-300 | head -20
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -n 1 gpurun_out/${TAG}_smoke.log | cut -c1-400
timeout 600 python bench.py > gpurun_out/${TAG}_bench.log 2>&1
echo "bench exit $?"; tail -n 1 gpurun_out/${TAG}_bench.log | cut -c1-1200
timeout 400 python bench.py --impl reference > gpurun_out/${TAG}_bench_ref.log 2>&1
echo "ref exit $?"; tail -n 1 gpurun_out/${TAG}_bench_ref.log | cut -c1-300
