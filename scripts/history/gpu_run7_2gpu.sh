#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 600 python -m pytest tests/test_multigpu.py -q -s -m gpu --maxfail=3 -p no:cacheprovider -k "multicast or auto" > gpurun_out/test_multicast.log 2>&1
echo "multicast test exit $?" >> gpurun_out/summary.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29721 \
   bench/sign_exchange_microbench.py --sizes 16,124 --skip-reference-above 0 --out gpurun_out/sign_exchange_w2_mc.json > gpurun_out/sign_exchange_w2_mc.log 2>&1
echo "sign_exchange exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -4 gpurun_out/test_multicast.log; grep '^{' gpurun_out/sign_exchange_w2_mc.log
