#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_gpu_model_kernels.py -x -q -p no:cacheprovider > gpurun_out/test_mk.log 2>&1; echo "model kernel tests exit $?" >> gpurun_out/summary.txt
timeout 300 python bench/model_kernels_microbench.py --out gpurun_out/model_kernels.json > gpurun_out/model_kernels.log 2>&1; echo "microbench exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -2 gpurun_out/test_mk.log; grep norm_ gpurun_out/model_kernels.log
