#!/bin/bash
# Optimizer-step timing with 10 back-to-back steps per sample, 2 GPUs (burst 1 = profiles/sign_exchange_w2_burst1_run30.json).
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
B=10
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 bench/sign_exchange_microbench.py --sizes 124,1000 --iters 8 --burst $B --skip-reference-above 0 --out gpurun_out/sign_exchange_w2_burst$B.json > gpurun_out/burst$B.log 2>&1
echo "burst $B exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; grep -h '^{' gpurun_out/burst10.log | cut -c1-400
