#!/bin/bash
# round-2 experiments: A/B of cheap launch-policy toggles on the bench line (each run: 20 steps, no CPU baseline)
TAG=${1:-r2x}
mkdir -p gpurun_out
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_$name.log 2>&1
  python - "$name" gpurun_out/${TAG}_$name.log <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print('%-22s ms/step %.4f  e2e %.4f  conv3x3 %.1f us  launches %d' % (sys.argv[1], d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['us'], d['gpu_launches']))
except Exception as e:
    print(sys.argv[1], 'FAILED', e)
PY
}
run base A=1
run base2 A=1
run balance NSG_CONV_BALANCE=1
run no_prebn_dbias NSG_PREBN_DBIAS=0
run both NSG_CONV_BALANCE=1 NSG_PREBN_DBIAS=0
run nopdl NSG_PDL=0
timeout 200 python -m pytest tests/test_gpu_tiled.py -m gpu -q -p no:cacheprovider > gpurun_out/${TAG}_tiled.log 2>&1; echo "tiled exit $?"; tail -n 3 gpurun_out/${TAG}_tiled.log | cut -c1-300
