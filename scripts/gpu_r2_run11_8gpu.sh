#!/bin/bash
# round 2, call 11 (8 GPUs): real-size fused tests, soak, GPT-2 / Llama-2-7B / DPO bench lines, sign-exchange microbench
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 8"
nvidia-smi topo -m > gpurun_out/r2_11_topology_8gpu.txt 2>&1
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q -k "real_size or small_grid or stalled or nvls or equals_split" > gpurun_out/r2_11_pytest_w8.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_11_pytest_w8.log
tail -4 gpurun_out/r2_11_pytest_w8.log
timeout 600 $TR --master-port 29531 scripts/soak_fused.py --launches 2000 --out gpurun_out/r2_11_soak_w8.json > gpurun_out/r2_11_soak_w8.log 2>&1; tail -1 gpurun_out/r2_11_soak_w8.log
timeout 600 $TR --master-port 29532 bench.py --gpus 8 --steps 6 --warmup 3 > gpurun_out/r2_11_bench_gpt2_n8.json 2> gpurun_out/r2_11_bench_gpt2_n8.err; cut -c1-200 gpurun_out/r2_11_bench_gpt2_n8.json
timeout 900 $TR --master-port 29533 bench/sign_exchange_microbench.py --sizes 124,1000,6740 --burst 1 --skip-reference-above 200 --out gpurun_out/r2_11_signx_w8_burst1.json > gpurun_out/r2_11_signx_w8_burst1.log 2>&1; tail -3 gpurun_out/r2_11_signx_w8_burst1.log | cut -c1-400
timeout 600 $TR --master-port 29534 bench/sign_exchange_microbench.py --sizes 1,16,124,1000 --burst 10 --skip-reference-above 0 --out gpurun_out/r2_11_signx_w8_burst10.json > gpurun_out/r2_11_signx_w8_burst10.log 2>&1; tail -2 gpurun_out/r2_11_signx_w8_burst10.log | cut -c1-400
timeout 900 $TR --master-port 29535 bench.py --gpus 8 --model llama-2-7b --batch 4 --ga 2 --steps 10 --warmup 3 > gpurun_out/r2_11_llama_ours_n8.json 2> gpurun_out/r2_11_llama_ours_n8.err; cut -c1-200 gpurun_out/r2_11_llama_ours_n8.json; tail -2 gpurun_out/r2_11_llama_ours_n8.err
timeout 900 $TR --master-port 29536 bench.py --gpus 8 --task dpo --steps 5 --warmup 3 > gpurun_out/r2_11_dpo_ours_n8.json 2> gpurun_out/r2_11_dpo_ours_n8.err; cut -c1-200 gpurun_out/r2_11_dpo_ours_n8.json; tail -2 gpurun_out/r2_11_dpo_ours_n8.err
