#!/bin/bash
# One GPU-box visit: tests, smoke, bench, then (only if the bench exited 0) the ncu launch list and one
# full capture of the kernel named by $NCU_KERNEL (regex).
# usage: tools/gpu_round.sh [tag] [bench args...]
TAG=${1:-r1}; shift
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q -p no:cacheprovider --timeout 300 > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/${TAG}_pytest.log | cut -c1-220
timeout 300 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -n 1 gpurun_out/${TAG}_smoke.log | cut -c1-300
timeout 600 python bench.py --steps 10 --warmup 3 "$@" > gpurun_out/${TAG}_bench.log 2>&1
BRC=$?; echo "bench exit $BRC"; tail -n 1 gpurun_out/${TAG}_bench.log | cut -c1-1500
if [ $BRC -eq 0 ]; then
  timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/${TAG}_plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv \
     --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/${TAG}_ncu.log 2>&1
  echo "ncu launches exit $?"
  if [ -n "$NCU_KERNEL" ]; then
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:$NCU_KERNEL -s ${NCU_SKIP:-6} -c 3 \
       -f -o gpurun_out/${TAG}_prof python bench.py --steps 2 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/${TAG}_ncufull.log 2>&1
    echo "ncu full exit $?"
  fi
fi
