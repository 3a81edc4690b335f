#!/bin/bash
# full ncu capture of the TMA conv kernels at one layer size: tools/ncu_ops.sh <layer index> <tag>
L=${1:-0}; TAG=${2:-ops}
mkdir -p gpurun_out
export BENCH_LAYER=$L BENCH_ITERS=2
python tools/bench_ops.py > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'conv_tma_kernel|wgrad_tma_kernel' -s 3 -c 6 -f \
    -o gpurun_out/${TAG}_prof python tools/bench_ops.py > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/${TAG}_ncu.log
