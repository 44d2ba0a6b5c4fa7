#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 400 python -m pytest tests/test_multigpu.py -q -s -m gpu --maxfail=3 -p no:cacheprovider -k "multicast or (average and float32)" > gpurun_out/test_multigpu8b.log 2>&1
echo "multigpu W=8 (multicast + rerun) exit $?" >> gpurun_out/summary.txt
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29731 \
   bench/sign_exchange_microbench.py --sizes 124,1000 --skip-reference-above 0 --out gpurun_out/sign_exchange_w8_mc.json > gpurun_out/sign_exchange_w8_mc.log 2>&1
echo "sign_exchange W=8 exit $?" >> gpurun_out/summary.txt
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29732 bench.py --gpus 8 --steps 4 --warmup 3 > gpurun_out/bench_ours_n8.log 2>&1
echo "bench ours n8 exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -3 gpurun_out/test_multigpu8b.log; grep '^{' gpurun_out/sign_exchange_w8_mc.log; tail -1 gpurun_out/bench_ours_n8.log | cut -c1-300
