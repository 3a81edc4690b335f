#!/bin/bash
# round-2 evidence visit (one B200): every GPU test file in its own process, smoke, bench (ours + reference arm),
# inference / VQ / op / audio timings, ncu launch list of the bench command, one full capture of the dominant kernels
TAG=${1:-r2c}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
: > gpurun_out/${TAG}_pytest.log
for f in tests/test_gpu_tma.py tests/test_gpu_ops.py tests/test_gpu_tma_model.py tests/test_gpu_model.py tests/test_gpu_main.py tests/test_gpu_data.py tests/test_gpu_images.py tests/test_gpu_audio.py tests/test_gpu_dist.py; do
  echo "=== $f" >> gpurun_out/${TAG}_pytest.log
  timeout 400 python -m pytest $f -m gpu -q -p no:cacheprovider -s >> gpurun_out/${TAG}_pytest.log 2>&1
  echo "pytest $f exit $?"
done
grep -E "^(FAILED|ERROR)|passed|failed|hang report\]" gpurun_out/${TAG}_pytest.log | cut -c1-260
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -n 2 gpurun_out/${TAG}_smoke.log | cut -c1-700
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.log 2>&1
echo "bench exit $?"; tail -n 1 gpurun_out/${TAG}_bench.log | cut -c1-3000
timeout 400 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/${TAG}_bench_ref.log 2>&1
echo "ref exit $?"; tail -n 1 gpurun_out/${TAG}_bench_ref.log | cut -c1-500
timeout 200 python tools/bench_infer.py > gpurun_out/${TAG}_infer.log 2>&1; echo "infer exit $?"; tail -n 1 gpurun_out/${TAG}_infer.log | cut -c1-1200
timeout 300 python tools/bench_vq.py --full > gpurun_out/${TAG}_vq.log 2>&1; echo "vq exit $?"; tail -n 25 gpurun_out/${TAG}_vq.log | cut -c1-200
timeout 200 python tools/bench_ops.py > gpurun_out/${TAG}_ops.log 2>&1; echo "ops exit $?"
timeout 200 python tools/bench_audio.py > gpurun_out/${TAG}_audio.log 2>&1; echo "audio exit $?"; tail -n 1 gpurun_out/${TAG}_audio.log | cut -c1-900
timeout 100 python tools/timeline_conv.py > gpurun_out/${TAG}_timeline.log 2>&1
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
   --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu launches exit $?"
timeout 600 ncu --set full --clock-control none --import-source on \
   -k regex:"conv_tma_kernel|wgrad_tma_kernel|vq_search_res|c1t_|scatter_a|bn_act_kernel|bn_bwd_kernel" -s 40 -c 48 \
   -f -o gpurun_out/${TAG}_prof python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncufull.log 2>&1
echo "ncu full exit $?"
timeout 300 ncu --set full --clock-control none -k regex:conv_tma_kernel -c 6 -f -o gpurun_out/${TAG}_roof \
   python bench.py --kernel-only > gpurun_out/${TAG}_roof.log 2>&1
echo "ncu roof exit $?"
