#!/bin/bash
# 1-GPU: verify fused linear, re-bench GPT-2, try the Llama-2-7B SFT and DPO configs, refresh launch list.
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/test_all_gpu.log 2>&1; echo "all gpu tests exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --steps 4 --warmup 3 > gpurun_out/bench_ours_n1.log 2>&1; echo "bench ours exit $?" >> gpurun_out/summary.txt
timeout 900 python bench.py --model llama-2-7b --batch 4 --ga 2 --steps 3 --warmup 3 > gpurun_out/bench_llama7b_n1.log 2>&1; echo "bench llama7b exit $?" >> gpurun_out/summary.txt
timeout 900 python -m dpo_llama2 --model_name_or_path llama-2-7b --use_lora False --synthetic_data --max_steps 4 --gradient_checkpointing False \
   --per_device_train_batch_size 4 --gradient_accumulation_steps 4 --lion --async_grad --lion_vote average --logging_steps 1 --save_steps 1000 --eval_steps 1000 \
   --output_dir /tmp/dpo7b > gpurun_out/dpo7b_n1.log 2>&1; echo "dpo 7b exit $?" >> gpurun_out/summary.txt
nvidia-smi --query-gpu=memory.used --format=csv >> gpurun_out/summary.txt
timeout 600 python bench.py --steps 1 --warmup 1 --ga 1 --no-e2e > gpurun_out/plain_small.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches2.csv \
   python bench.py --steps 1 --warmup 1 --ga 1 --no-e2e > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -4 gpurun_out/test_all_gpu.log; tail -1 gpurun_out/bench_ours_n1.log | cut -c1-330; tail -2 gpurun_out/bench_llama7b_n1.log | cut -c1-600; tail -12 gpurun_out/dpo7b_n1.log
