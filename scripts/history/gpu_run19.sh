#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/test_all_gpu.log 2>&1; echo "all gpu tests exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --no-e2e > gpurun_out/bench_ours_n1.log 2>&1; echo "bench exit $?" >> gpurun_out/summary.txt
DLION_NORM_CTA_PER_ROW=1 timeout 600 python bench.py --no-e2e > gpurun_out/bench_cta_norm.log 2>&1; echo "bench cta-norm exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --steps 1 --warmup 1 --ga 1 --no-e2e > gpurun_out/plain_small.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches4.csv \
   python bench.py --steps 1 --warmup 1 --ga 1 --no-e2e > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -3 gpurun_out/test_all_gpu.log; tail -1 gpurun_out/bench_ours_n1.log | cut -c1-260; tail -1 gpurun_out/bench_cta_norm.log | cut -c1-260
