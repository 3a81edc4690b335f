#!/bin/bash
# eight-GPU visit (charged 8x): ONE bench run in synchronised-BatchNorm mode; the line carries the local-BN mode too
TAG=${1:-r2g8}
N=${2:-8}
mkdir -p gpurun_out
NSG_PEER_IN_KERNEL=${PEER_IN_KERNEL:-1} NSG_BENCH_SYNC_BN=1 NCCL_DEBUG=WARN timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 \
   bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${TAG}_bench${N}_sb1.log 2>&1
echo "bench N=$N sync_bn=1 exit $?"; grep '"metric"' gpurun_out/${TAG}_bench${N}_sb1.log | tail -n 1 | cut -c1-2400
grep -h "hang report\|peer-memory\|Error\|error" gpurun_out/${TAG}_bench*.log | head -10
