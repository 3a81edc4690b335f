#!/bin/bash
# A/B of the convergent-warp (elect.sync) MMA issue loop in conv_tma_kernel (NSG_CONV_ELECT)
TAG=${1:-r2u}
mkdir -p gpurun_out
NSG_CONV_ELECT=1 timeout 600 python -m pytest tests/test_gpu_tma.py tests/test_gpu_tma_model.py -x -m gpu -q -p no:cacheprovider > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest (elect) exit $?"; tail -n 3 gpurun_out/${TAG}_pytest.log | cut -c1-300; grep -E "^(FAILED|ERROR)|hang report\]|^E  " gpurun_out/${TAG}_pytest.log | head -10 | cut -c1-300
for n in 1 0 1 0; do
  NSG_CONV_ELECT=$n timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_bench_e$n.log 2>&1
  echo "elect=$n bench exit $?"; tail -n 1 gpurun_out/${TAG}_bench_e$n.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['us'], d['roofline']['frac'])" 2>&1 | tail -n 1 | cut -c1-300
done
for n in 1 0; do
  NSG_CONV_ELECT=$n timeout 200 python tools/bench_ops.py > gpurun_out/${TAG}_ops_e$n.log 2>&1; echo "elect=$n ops exit $?"
  grep -E "fwd\+stats|dgrad\+mask" gpurun_out/${TAG}_ops_e$n.log | grep "^conv" | cut -c1-100
  NSG_CONV_ELECT=$n timeout 120 python tools/timeline_conv.py > gpurun_out/${TAG}_timeline_e$n.log 2>&1; echo "timeline exit $?"; grep -E "^(3x3|1x1|4x4)" gpurun_out/${TAG}_timeline_e$n.log | cut -c1-420
done
