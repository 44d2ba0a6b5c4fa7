#!/bin/bash
# round 2, call 6 (2 GPUs): all tests, soak, GPT-2 N=2, Llama-2-7B ours N=1/2 + reference N=1/2, DPO ours N=2
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_6_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_6_pytest.log
tail -6 gpurun_out/r2_6_pytest.log
timeout 600 $TR --nproc-per-node 2 --master-port 29521 scripts/soak_fused.py --launches 2000 --out gpurun_out/r2_6_soak_w2.json > gpurun_out/r2_6_soak_w2.log 2>&1; tail -2 gpurun_out/r2_6_soak_w2.log
timeout 600 $TR --nproc-per-node 2 --master-port 29522 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/r2_6_bench_gpt2_n2.json 2> gpurun_out/r2_6_bench_gpt2_n2.err; cat gpurun_out/r2_6_bench_gpt2_n2.json
timeout 900 python bench.py --model llama-2-7b --batch 4 --ga 2 --steps 10 --warmup 3 > gpurun_out/r2_6_llama_ours_n1.json 2> gpurun_out/r2_6_llama_ours_n1.err; cat gpurun_out/r2_6_llama_ours_n1.json; tail -2 gpurun_out/r2_6_llama_ours_n1.err
timeout 900 $TR --nproc-per-node 2 --master-port 29523 bench.py --gpus 2 --model llama-2-7b --batch 4 --ga 2 --steps 10 --warmup 3 > gpurun_out/r2_6_llama_ours_n2.json 2> gpurun_out/r2_6_llama_ours_n2.err; cat gpurun_out/r2_6_llama_ours_n2.json; tail -2 gpurun_out/r2_6_llama_ours_n2.err
timeout 900 python bench.py --impl reference --model llama-2-7b --batch 4 --ga 2 --steps 6 --warmup 3 > gpurun_out/r2_6_llama_ref_n1.json 2> gpurun_out/r2_6_llama_ref_n1.err; cat gpurun_out/r2_6_llama_ref_n1.json; tail -2 gpurun_out/r2_6_llama_ref_n1.err
timeout 1200 $TR --nproc-per-node 2 --master-port 29524 bench.py --impl reference --gpus 2 --model llama-2-7b --batch 4 --ga 2 --steps 2 --warmup 1 --no-e2e > gpurun_out/r2_6_llama_ref_n2.json 2> gpurun_out/r2_6_llama_ref_n2.err; cat gpurun_out/r2_6_llama_ref_n2.json; tail -2 gpurun_out/r2_6_llama_ref_n2.err
timeout 900 $TR --nproc-per-node 2 --master-port 29525 bench.py --gpus 2 --task dpo --steps 6 --warmup 3 > gpurun_out/r2_6_dpo_ours_n2.json 2> gpurun_out/r2_6_dpo_ours_n2.err; cat gpurun_out/r2_6_dpo_ours_n2.json; tail -2 gpurun_out/r2_6_dpo_ours_n2.err
