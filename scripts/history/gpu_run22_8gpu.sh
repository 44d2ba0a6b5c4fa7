#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29761 \
   bench/sign_exchange_microbench.py --sizes 6738 --iters 5 --skip-reference-above 0 --out gpurun_out/sign_exchange_w8_7b.json > gpurun_out/sign_exchange_w8_7b.log 2>&1
echo "sign_exchange 7B exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; grep '^{' gpurun_out/sign_exchange_w8_7b.log; grep -E "OutOfMemory" gpurun_out/sign_exchange_w8_7b.log | head -2
