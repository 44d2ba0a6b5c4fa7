#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 600 python bench.py --model llama-2-7b --batch 4 --ga 2 --steps 3 --warmup 3 --lora > gpurun_out/bench_llama7b_lora_bf16.log 2>&1; echo "lora bf16 exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --model llama-2-7b --batch 4 --ga 2 --steps 3 --warmup 3 --lora --load-in-4bit > gpurun_out/bench_llama7b_lora_nf4.log 2>&1; echo "lora nf4 exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
for f in gpurun_out/bench_llama7b_lora_bf16.log gpurun_out/bench_llama7b_lora_nf4.log; do tail -1 $f | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['ms_per_step'],1), d['clocks']['sm_mhz'], d.get('peak_mem_gb'), d['e2e']['value'])"; done
