#!/bin/bash
# Reference arm on the Llama-2-7B SFT config (per-device batch 4x1024, GA 2), 1 GPU; ours for the same box.
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 900 python bench.py --impl reference --model llama-2-7b --batch 4 --ga 2 --steps 3 --warmup 3 > gpurun_out/bench_ref_llama7b.log 2>&1; echo "reference llama exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --model llama-2-7b --batch 4 --ga 2 --steps 3 --warmup 3 > gpurun_out/bench_ours_llama7b.log 2>&1; echo "ours llama exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
for f in gpurun_out/bench_ref_llama7b.log gpurun_out/bench_ours_llama7b.log; do tail -1 $f | cut -c1-250; done
