#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_gpu_model_kernels.py tests/test_gpu_lt_gemm.py -x -q -p no:cacheprovider > gpurun_out/test_mk.log 2>&1; echo "model kernel tests exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py > gpurun_out/bench_ours_n1.log 2>&1; echo "bench exit $?" >> gpurun_out/summary.txt
DLION_FUSE_COLSUM=0 timeout 600 python bench.py --no-e2e > gpurun_out/bench_nofuse.log 2>&1; echo "bench (two-launch colsum) exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -3 gpurun_out/test_mk.log; tail -1 gpurun_out/bench_ours_n1.log | cut -c1-260; tail -1 gpurun_out/bench_nofuse.log | cut -c1-260
