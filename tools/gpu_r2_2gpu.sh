#!/bin/bash
# two-GPU visit: the data-parallel equivalence tests (sync-BN over peer memory, local BN), then the bench at N = 2 in both
# BatchNorm modes (each line carries the other mode as a sub-object), with and without the in-kernel statistic exchange
TAG=${1:-r2g2}
N=${2:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/${TAG}_topo.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_dist.py -m gpu -q -p no:cacheprovider -s > gpurun_out/${TAG}_dist.log 2>&1
echo "dist pytest exit $?"; grep -E "^(FAILED|ERROR)|passed|failed|hang report\]|peer-memory|sync_bn|rel" gpurun_out/${TAG}_dist.log | cut -c1-260 | head -30
for sb in 1 0; do
  NSG_BENCH_SYNC_BN=$sb timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
     bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${TAG}_bench${N}_sb$sb.log 2>&1
  echo "bench N=$N sync_bn=$sb exit $?"; grep '"metric"' gpurun_out/${TAG}_bench${N}_sb$sb.log | tail -n 1 | cut -c1-2200
done
NSG_PEER_SYNC=0 NSG_BENCH_SYNC_BN=1 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
   bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${TAG}_bench${N}_sb1_nccl.log 2>&1
echo "bench N=$N sync_bn=1 (NCCL statistics) exit $?"; grep '"metric"' gpurun_out/${TAG}_bench${N}_sb1_nccl.log | tail -n 1 | cut -c1-1500
grep -h "hang report\|peer-memory\|Error\|error" gpurun_out/${TAG}_bench*.log | head -20
