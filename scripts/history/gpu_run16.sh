#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/test_all_gpu.log 2>&1; echo "all gpu tests exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py > gpurun_out/bench_ours_n1.log 2>&1; echo "bench exit $?" >> gpurun_out/summary.txt
timeout 900 python bench.py --model llama-2-7b --batch 4 --ga 2 --steps 3 --warmup 3 --no-e2e > gpurun_out/bench_llama7b_n1.log 2>&1; echo "bench llama7b exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -3 gpurun_out/test_all_gpu.log; tail -1 gpurun_out/bench_ours_n1.log | cut -c1-260;  tail -1 gpurun_out/bench_llama7b_n1.log | cut -c1-260
