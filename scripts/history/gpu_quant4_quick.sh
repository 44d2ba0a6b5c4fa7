#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 600 python -m pytest tests/test_gpu_quant4.py -x -q -p no:cacheprovider > gpurun_out/test_quant4.log 2>&1; echo "quant4 tests exit $?" >> gpurun_out/summary.txt
timeout 300 python bench/model_kernels_microbench.py --out gpurun_out/model_kernels.json > gpurun_out/model_kernels.log 2>&1; echo "microbench exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -3 gpurun_out/test_quant4.log; tail -4 gpurun_out/model_kernels.log
