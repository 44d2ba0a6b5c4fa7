#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_gpu_model_kernels.py -x -q -m gpu -p no:cacheprovider > gpurun_out/test_model.log 2>&1; echo "model kernel tests exit $?" >> gpurun_out/summary.txt
timeout 300 python bench/model_kernels_microbench.py > gpurun_out/mk.log 2>&1; echo "mk exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --no-e2e > gpurun_out/bench_ours_n1.log 2>&1; echo "bench exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -2 gpurun_out/test_model.log; grep ce_fwd gpurun_out/mk.log; tail -1 gpurun_out/bench_ours_n1.log | cut -c1-260
