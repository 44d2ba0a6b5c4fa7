#!/bin/bash
# round 2, call 12 (4 GPUs): GPT-2 and Llama-2-7B bench lines, sign-exchange microbench at W = 4
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 4"
timeout 600 $TR --master-port 29541 bench.py --gpus 4 --steps 6 --warmup 3 > gpurun_out/r2_12_bench_gpt2_n4.json 2> gpurun_out/r2_12_bench_gpt2_n4.err; cut -c1-200 gpurun_out/r2_12_bench_gpt2_n4.json
timeout 900 $TR --master-port 29542 bench.py --gpus 4 --model llama-2-7b --batch 4 --ga 2 --steps 10 --warmup 3 > gpurun_out/r2_12_llama_ours_n4.json 2> gpurun_out/r2_12_llama_ours_n4.err; cut -c1-200 gpurun_out/r2_12_llama_ours_n4.json; tail -2 gpurun_out/r2_12_llama_ours_n4.err
timeout 600 $TR --master-port 29543 bench/sign_exchange_microbench.py --sizes 124,1000,6740 --burst 1 --skip-reference-above 0 --out gpurun_out/r2_12_signx_w4_burst1.json > gpurun_out/r2_12_signx_w4_burst1.log 2>&1; tail -2 gpurun_out/r2_12_signx_w4_burst1.log | cut -c1-300
timeout 600 $TR --master-port 29544 bench/sign_exchange_microbench.py --sizes 1,16,124,1000 --burst 10 --skip-reference-above 0 --out gpurun_out/r2_12_signx_w4_burst10.json > gpurun_out/r2_12_signx_w4_burst10.log 2>&1; tail -2 gpurun_out/r2_12_signx_w4_burst10.log | cut -c1-300
