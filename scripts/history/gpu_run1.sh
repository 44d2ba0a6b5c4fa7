#!/bin/bash
# First GPU pass: kernel numerics, bench (ours + reference), launch list, microbench.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
for f in test_gpu_lion_kernels test_gpu_model_kernels test_gpu_fused; do
  timeout 900 python -m pytest tests/$f.py -q -s -m gpu --maxfail=40 -p no:cacheprovider > gpurun_out/$f.log 2>&1
  echo "$f exit $?" >> gpurun_out/summary.txt
done
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/summary.txt
timeout 600 python bench/optimizer_microbench.py --out gpurun_out/microbench.json > gpurun_out/microbench.log 2>&1; echo "microbench exit $?" >> gpurun_out/summary.txt
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_ours.log 2>&1; echo "bench ours exit $?" >> gpurun_out/summary.txt
timeout 900 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/bench_ref.log 2>&1; echo "bench ref exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --steps 1 --warmup 1 --ga 1 --no-e2e > gpurun_out/plain_small.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches.csv \
   python bench.py --steps 1 --warmup 1 --ga 1 --no-e2e > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
tail -5 gpurun_out/bench_ours.log gpurun_out/bench_ref.log
