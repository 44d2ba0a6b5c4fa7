#!/bin/bash
# 8-GPU pass: real W=8 fused exchange, microbench, headline bench both arms, Llama-2-7B SFT and DPO configs.
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
nvidia-smi topo -m > gpurun_out/topo8.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_multigpu.py -q -s -m gpu --maxfail=2 -p no:cacheprovider > gpurun_out/test_multigpu8.log 2>&1
echo "test_multigpu W=8 exit $?" >> gpurun_out/summary.txt
timeout 900 $TR --master-port 29711 bench/sign_exchange_microbench.py --sizes 1,16,124,1000 --out gpurun_out/sign_exchange_w8.json > gpurun_out/sign_exchange_w8.log 2>&1
echo "sign_exchange W=8 exit $?" >> gpurun_out/summary.txt
timeout 900 $TR --master-port 29712 bench.py --gpus 8 --steps 4 --warmup 3 > gpurun_out/bench_ours_n8.log 2>&1; echo "bench ours n8 exit $?" >> gpurun_out/summary.txt
timeout 900 $TR --master-port 29713 bench.py --gpus 8 --steps 3 --warmup 3 --impl reference > gpurun_out/bench_ref_n8.log 2>&1; echo "bench ref n8 exit $?" >> gpurun_out/summary.txt
timeout 900 $TR --master-port 29714 bench.py --gpus 8 --model llama-2-7b --batch 4 --ga 2 --steps 3 --warmup 3 > gpurun_out/bench_llama7b_n8.log 2>&1; echo "bench llama7b n8 exit $?" >> gpurun_out/summary.txt
timeout 900 $TR --master-port 29715 -m dpo_llama2 --model_name_or_path llama-2-7b --use_lora False --synthetic_data --max_steps 6 --gradient_checkpointing False \
   --per_device_train_batch_size 4 --gradient_accumulation_steps 4 --lion --async_grad --lion_vote average --lion_drop_prob 0.2 --logging_steps 1 --save_steps 1000 --eval_steps 1000 \
   --output_dir /tmp/dpo7b > gpurun_out/dpo7b_n8.log 2>&1; echo "dpo 7b n8 exit $?" >> gpurun_out/summary.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29716 bench.py --gpus 4 --steps 4 --warmup 3 > gpurun_out/bench_ours_n4.log 2>&1; echo "bench ours n4 exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -3 gpurun_out/test_multigpu8.log; grep '^{' gpurun_out/sign_exchange_w8.log; for f in bench_ours_n8 bench_ref_n8 bench_llama7b_n8 bench_ours_n4; do tail -1 gpurun_out/$f.log | cut -c1-420; done; grep -E "live_workers|train_tokens_per_second" gpurun_out/dpo7b_n8.log | tail -4 | cut -c1-400
