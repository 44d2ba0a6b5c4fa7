#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29751 bench.py --gpus 8 > gpurun_out/bench_ours_n8.log 2>&1; echo "bench ours n8 exit $?" >> gpurun_out/summary.txt
timeout 400 $TR --master-port 29752 bench/sign_exchange_microbench.py --sizes 6738 --iters 5 --skip-reference-above 0 --out gpurun_out/sign_exchange_w8_7b.json > gpurun_out/sign_exchange_w8_7b.log 2>&1
echo "sign_exchange 7B exit $?" >> gpurun_out/summary.txt
timeout 400 $TR --master-port 29753 bench.py --gpus 8 --model llama-2-7b --batch 4 --ga 2 --steps 3 --warmup 3 > gpurun_out/bench_llama7b_n8.log 2>&1; echo "bench llama7b n8 exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -1 gpurun_out/bench_ours_n8.log | cut -c1-300; grep '^{' gpurun_out/sign_exchange_w8_7b.log; tail -3 gpurun_out/sign_exchange_w8_7b.log | cut -c1-300; tail -1 gpurun_out/bench_llama7b_n8.log | cut -c1-300
