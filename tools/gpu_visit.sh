#!/bin/bash
# One GPU-box visit: tests, smoke, bench in the three modes, then the ncu launch list of the bf16 step and full
# captures of the kernels named by $NCU_KERNEL (regex).    usage: tools/gpu_visit.sh [tag]
TAG=${1:-r1v}; shift
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q -p no:cacheprovider --timeout 300 > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/${TAG}_pytest.log | cut -c1-220
timeout 300 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -n 1 gpurun_out/${TAG}_smoke.log | cut -c1-300
for MODE in ${MODES:-fp32 tc32 bf16 tma}; do
  EXTRA="--no-cpu-baseline"; [ $MODE = tma ] && EXTRA=""
  timeout 600 python bench.py --steps 20 --warmup 5 --mode $MODE $EXTRA > gpurun_out/${TAG}_bench_$MODE.log 2>&1
  echo "bench $MODE exit $?"; tail -n 1 gpurun_out/${TAG}_bench_$MODE.log | cut -c1-1800
done
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.log 2>&1; echo "ref exit $?"; tail -n 1 gpurun_out/${TAG}_bench_ref.log | cut -c1-600
M=${NCU_MODE:-tma}
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --mode $M > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
   --log-file gpurun_out/${TAG}_launches_$M.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --mode $M > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu launches exit $?"
if [ -n "$NCU_KERNEL" ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$NCU_KERNEL -s ${NCU_SKIP:-20} -c ${NCU_COUNT:-12} \
     -f -o gpurun_out/${TAG}_prof_$M python bench.py --steps 2 --warmup 3 --no-cpu-baseline --mode $M > gpurun_out/${TAG}_ncufull.log 2>&1
  echo "ncu full exit $?"
fi
