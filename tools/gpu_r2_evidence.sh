#!/bin/bash
# round-2 closing evidence (one B200): the GPU suite in ONE process with -x (as the driver runs it), smoke(), both bench
# arms, inference / op / VQ timings, the ncu launch list of the bench command, one full capture of the dominant kernel
TAG=${1:-r2z}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
timeout 1200 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest (one process, -x) exit $?"; tail -n 4 gpurun_out/${TAG}_pytest.log | cut -c1-300
grep -E "^(FAILED|ERROR)|hang report\]|^E  " gpurun_out/${TAG}_pytest.log | cut -c1-300 | head -20
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -n 1 gpurun_out/${TAG}_smoke.log | cut -c1-400
timeout 600 python bench.py > gpurun_out/${TAG}_bench.log 2>&1
echo "bench exit $?"; tail -n 1 gpurun_out/${TAG}_bench.log | cut -c1-2500
timeout 400 python bench.py --impl reference > gpurun_out/${TAG}_bench_ref.log 2>&1
echo "ref exit $?"; tail -n 1 gpurun_out/${TAG}_bench_ref.log | cut -c1-500
timeout 200 python tools/bench_infer.py > gpurun_out/${TAG}_infer.log 2>&1; echo "infer exit $?"; tail -n 1 gpurun_out/${TAG}_infer.log | cut -c1-600
timeout 200 python tools/bench_ops.py > gpurun_out/${TAG}_ops.log 2>&1; echo "ops exit $?"
timeout 300 python tools/bench_vq.py > gpurun_out/${TAG}_vq.log 2>&1; echo "vq exit $?"; grep "vq_search" gpurun_out/${TAG}_vq.log | grep "mode=1" | cut -c1-160
timeout 100 python tools/timeline_conv.py > gpurun_out/${TAG}_timeline.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
   --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu launches exit $?"
python tools/launch_summary.py gpurun_out/${TAG}_launches.csv > gpurun_out/${TAG}_launch_shares.txt 2>&1; head -n 12 gpurun_out/${TAG}_launch_shares.txt | cut -c1-140
timeout 300 ncu --set full --clock-control none --import-source on -k regex:conv_tma_kernel -c 6 -f -o /tmp/${TAG}_roof \
   python bench.py --kernel-only > gpurun_out/${TAG}_roof.log 2>&1
echo "ncu roof exit $?"
ncu -i /tmp/${TAG}_roof.ncu-rep --page raw --csv > gpurun_out/${TAG}_roof_raw.csv 2>/dev/null
ncu -i /tmp/${TAG}_roof.ncu-rep --page details > gpurun_out/${TAG}_roof_details.txt 2>/dev/null
python tools/ncu_traffic.py gpurun_out/${TAG}_roof_raw.csv --out gpurun_out/${TAG}_roofline_traffic.json 2>&1 | tail -n 2
ls -la gpurun_out | head -40
