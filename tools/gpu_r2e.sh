#!/bin/bash
# round-2 visit: A/B of the BatchNorm fusion (NSG_FUSE_BN), launch lists of both, a SMALL full capture exported to csv on
# the box (the .ncu-rep files stay there: gpurun_out is capped at 64 MiB), N4 tests, details of the failing tests
TAG=${1:-r2e}
mkdir -p gpurun_out
: > gpurun_out/${TAG}_pytest.log
for f in tests/test_gpu_n4.py tests/test_gpu_tma_model.py tests/test_gpu_images.py; do
  echo "=== $f" >> gpurun_out/${TAG}_pytest.log
  timeout 400 python -m pytest $f -m gpu -q -p no:cacheprovider -s >> gpurun_out/${TAG}_pytest.log 2>&1
  echo "pytest $f exit $?"
done
grep -E "^(FAILED|ERROR)|passed|failed|hang report\]" gpurun_out/${TAG}_pytest.log | cut -c1-260
for fb in 1 0; do
  NSG_FUSE_BN=$fb timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_bench_fb$fb.log 2>&1
  echo "bench fuse_bn=$fb exit $?"; tail -n 1 gpurun_out/${TAG}_bench_fb$fb.log | cut -c1-1400
  NSG_FUSE_BN=$fb timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv \
     --log-file gpurun_out/${TAG}_launches_fb$fb.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_fb$fb.log 2>&1
  echo "ncu launches fuse_bn=$fb exit $?"
  python tools/launch_summary.py gpurun_out/${TAG}_launches_fb$fb.csv > gpurun_out/${TAG}_launch_shares_fb$fb.txt 2>&1
  head -n 30 gpurun_out/${TAG}_launch_shares_fb$fb.txt | cut -c1-160
done
timeout 600 ncu --set full --clock-control none \
   -k regex:"conv_tma_kernel|wgrad_tma_kernel|vq_search_res|c1t_|scatter_a" -s 60 -c 36 \
   -f -o /tmp/${TAG}_prof python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncufull.log 2>&1
echo "ncu full exit $?"
ncu -i /tmp/${TAG}_prof.ncu-rep --page raw --csv > gpurun_out/${TAG}_prof_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/${TAG}_prof_raw.csv > gpurun_out/${TAG}_ncu_full_summary.txt 2>&1
timeout 300 ncu --set full --clock-control none -k regex:conv_tma_kernel -c 6 -f -o /tmp/${TAG}_roof \
   python bench.py --kernel-only > gpurun_out/${TAG}_roof.log 2>&1
echo "ncu roof exit $?"
ncu -i /tmp/${TAG}_roof.ncu-rep --page raw --csv > gpurun_out/${TAG}_roof_raw.csv 2>/dev/null
timeout 300 ncu --set full --clock-control none -k regex:vq_search -c 4 -f -o /tmp/${TAG}_vq \
   python tools/vq_one.py > gpurun_out/${TAG}_vqone.log 2>&1
echo "ncu vq exit $?"
ncu -i /tmp/${TAG}_vq.ncu-rep --page raw --csv > gpurun_out/${TAG}_vq_raw.csv 2>/dev/null
ncu -i /tmp/${TAG}_vq.ncu-rep --page details > gpurun_out/${TAG}_vq_details.txt 2>/dev/null
du -sh gpurun_out
