#!/bin/bash
# Round-end evidence visit (one B200): tests, smoke, the bench line and its reference arm, inference latency, VQ kernels,
# then the ncu launch list of the same bench command and one full capture of the dominant kernels.
#   usage: tools/gpu_final.sh [tag]        outputs under gpurun_out/<tag>_*
TAG=${1:-r1f}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
timeout 600 python -m pytest tests -m gpu -q -p no:cacheprovider --timeout 200 > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/${TAG}_pytest.log | cut -c1-220
timeout 200 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -n 1 gpurun_out/${TAG}_smoke.log | cut -c1-300
timeout 400 python bench.py > gpurun_out/${TAG}_bench.log 2>&1; echo "bench exit $?"; tail -n 1 gpurun_out/${TAG}_bench.log | cut -c1-2500
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.log 2>&1; echo "ref exit $?"; tail -n 1 gpurun_out/${TAG}_bench_ref.log | cut -c1-400
timeout 200 python tools/bench_infer.py > gpurun_out/${TAG}_infer.log 2>&1; echo "infer exit $?"; tail -n 1 gpurun_out/${TAG}_infer.log | cut -c1-1200
timeout 200 python tools/bench_vq.py --full > gpurun_out/${TAG}_vq.log 2>&1; echo "vq exit $?"
timeout 200 python tools/bench_ops.py > gpurun_out/${TAG}_ops.log 2>&1; echo "ops exit $?"
timeout 100 python tools/timeline_conv.py > gpurun_out/${TAG}_timeline.log 2>&1; timeout 100 python tools/timeline_wgrad.py > gpurun_out/${TAG}_timeline_wgrad.log 2>&1
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
   --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu launches exit $?"
timeout 600 ncu --set full --clock-control none --import-source on \
   -k regex:"conv_tma_kernel|wgrad_tma_kernel|vq_search_res|c1t_|scatter_add_kernel|bn_act_kernel|bn_bwd_kernel" -s 40 -c 40 \
   -f -o gpurun_out/${TAG}_prof python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncufull.log 2>&1
echo "ncu full exit $?"
