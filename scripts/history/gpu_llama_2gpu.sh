#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29547 bench.py --gpus 2 --model llama-2-7b --batch 4 --ga 2 --steps 3 --warmup 3 > gpurun_out/bench_ours_llama7b_n2.log 2>&1
echo "llama n2 exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -1 gpurun_out/bench_ours_llama7b_n2.log | cut -c1-330
