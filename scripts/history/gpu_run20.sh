#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_gpu_model_kernels.py -x -q -m gpu -p no:cacheprovider > gpurun_out/test_model.log 2>&1; echo "model tests exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py > gpurun_out/bench_ours_n1.log 2>&1; echo "bench exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --steps 1 --warmup 1 --ga 1 --no-e2e > gpurun_out/plain_small.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches5.csv \
   python bench.py --steps 1 --warmup 1 --ga 1 --no-e2e > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -2 gpurun_out/test_model.log; tail -1 gpurun_out/bench_ours_n1.log | cut -c1-260
