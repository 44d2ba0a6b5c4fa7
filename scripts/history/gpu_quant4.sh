#!/bin/bash
# 4-bit codec: GPU tests, the model-kernel microbenchmark (includes the q4 rows), then Llama-2-7B
# LoRA SFT steps (per-device batch 4x1024, GA 2) with a bf16 base and with an NF4 base.
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 600 python -m pytest tests/test_gpu_quant4.py -x -q -p no:cacheprovider > gpurun_out/test_quant4.log 2>&1; echo "quant4 tests exit $?" >> gpurun_out/summary.txt
timeout 300 python bench/model_kernels_microbench.py --out gpurun_out/model_kernels.json > gpurun_out/model_kernels.log 2>&1; echo "microbench exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --model llama-2-7b --batch 4 --ga 2 --steps 3 --warmup 3 --lora > gpurun_out/bench_llama7b_lora_bf16.log 2>&1; echo "lora bf16 exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --model llama-2-7b --batch 4 --ga 2 --steps 3 --warmup 3 --lora --load-in-4bit > gpurun_out/bench_llama7b_lora_nf4.log 2>&1; echo "lora nf4 exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -3 gpurun_out/test_quant4.log; tail -4 gpurun_out/model_kernels.log
tail -1 gpurun_out/bench_llama7b_lora_bf16.log | cut -c1-700; tail -1 gpurun_out/bench_llama7b_lora_nf4.log | cut -c1-700
