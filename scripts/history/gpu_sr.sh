#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_gpu_stochastic_rounding.py tests/test_gpu_fused.py tests/test_gpu_lion_kernels.py -x -q -p no:cacheprovider > gpurun_out/test_sr.log 2>&1; echo "sr + optimizer kernel tests exit $?" >> gpurun_out/summary.txt
timeout 600 python bench/optimizer_microbench.py > gpurun_out/opt_microbench.log 2>&1; echo "optimizer microbench exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -4 gpurun_out/test_sr.log; tail -25 gpurun_out/opt_microbench.log
