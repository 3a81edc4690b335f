#!/bin/bash
# two-GPU check after a scheduling change: the data-parallel equivalence tests and one bench line per BatchNorm mode
TAG=${1:-r2g2s}
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_dist.py -m gpu -q -p no:cacheprovider > gpurun_out/${TAG}_dist.log 2>&1
echo "dist pytest exit $?"; tail -n 2 gpurun_out/${TAG}_dist.log | cut -c1-200
NSG_BENCH_SYNC_BN=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 \
   bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/${TAG}_bench2_sb0.log 2>&1
echo "bench N=2 exit $?"; grep '"metric"' gpurun_out/${TAG}_bench2_sb0.log | tail -n 1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['e2e']['value'], d.get('sync_bn',{}).get('ms_per_step'), d.get('sync_bn',{}).get('value'))"
