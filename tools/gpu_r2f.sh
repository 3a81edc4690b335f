#!/bin/bash
# round-2 visit: whole suite (one process per file), smoke, the bench line + reference arm, launch list, ops timings
TAG=${1:-r2f}
mkdir -p gpurun_out
: > gpurun_out/${TAG}_pytest.log
for f in tests/test_gpu_tma.py tests/test_gpu_ops.py tests/test_gpu_tma_model.py tests/test_gpu_model.py tests/test_gpu_main.py tests/test_gpu_data.py tests/test_gpu_images.py tests/test_gpu_audio.py tests/test_gpu_n4.py tests/test_gpu_dist.py; do
  echo "=== $f" >> gpurun_out/${TAG}_pytest.log
  timeout 400 python -m pytest $f -m gpu -q -p no:cacheprovider -s >> gpurun_out/${TAG}_pytest.log 2>&1
  echo "pytest $f exit $?"
done
grep -E "^(FAILED|ERROR)|passed|failed|hang report\]" gpurun_out/${TAG}_pytest.log | cut -c1-260
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -n 1 gpurun_out/${TAG}_smoke.log | cut -c1-700
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.log 2>&1
echo "bench exit $?"; tail -n 1 gpurun_out/${TAG}_bench.log | cut -c1-3000
timeout 400 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/${TAG}_bench_ref.log 2>&1
echo "ref exit $?"; tail -n 1 gpurun_out/${TAG}_bench_ref.log | cut -c1-300
timeout 200 python tools/bench_infer.py > gpurun_out/${TAG}_infer.log 2>&1; echo "infer exit $?"; tail -n 1 gpurun_out/${TAG}_infer.log | cut -c1-900
timeout 200 python tools/bench_ops.py > gpurun_out/${TAG}_ops.log 2>&1; echo "ops exit $?"; tail -n 30 gpurun_out/${TAG}_ops.log | cut -c1-200
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv \
   --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu launches exit $?"
python tools/launch_summary.py gpurun_out/${TAG}_launches.csv > gpurun_out/${TAG}_launch_shares.txt 2>&1
head -n 24 gpurun_out/${TAG}_launch_shares.txt | cut -c1-160
timeout 600 ncu --set full --clock-control none \
   -k regex:"conv_tma_kernel|wgrad_tma_kernel|vq_search_res|c1t_|scatter_a|bn_act_kernel|bn_bwd_kernel" -s 80 -c 44 \
   -f -o /tmp/${TAG}_prof python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncufull.log 2>&1
echo "ncu full exit $?"
ncu -i /tmp/${TAG}_prof.ncu-rep --page raw --csv > gpurun_out/${TAG}_prof_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/${TAG}_prof_raw.csv > gpurun_out/${TAG}_ncu_full_summary.txt 2>&1
timeout 300 ncu --set full --clock-control none -k regex:conv_tma_kernel -c 6 -f -o /tmp/${TAG}_roof \
   python bench.py --kernel-only > gpurun_out/${TAG}_roof.log 2>&1
echo "ncu roof exit $?"
ncu -i /tmp/${TAG}_roof.ncu-rep --page raw --csv > gpurun_out/${TAG}_roof_raw.csv 2>/dev/null
du -sh gpurun_out
