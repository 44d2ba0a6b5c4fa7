#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 4 --steps 4 --warmup 3 > gpurun_out/bench_ours_n4.log 2>&1
echo "bench n4 exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -1 gpurun_out/bench_ours_n4.log | cut -c1-330
