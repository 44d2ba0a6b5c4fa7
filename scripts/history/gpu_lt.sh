#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 600 python -m pytest tests/test_gpu_lt_gemm.py tests/test_gpu_model_kernels.py -x -q -p no:cacheprovider > gpurun_out/test_lt.log 2>&1; echo "lt tests exit $?" >> gpurun_out/summary.txt
timeout 600 python bench/lt_epilogue_microbench.py --out gpurun_out/lt_epilogue.json > gpurun_out/lt_epilogue.log 2>&1; echo "lt bench exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py > gpurun_out/bench_ours_n1.log 2>&1; echo "bench exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -4 gpurun_out/test_lt.log; grep -E "^(unfused|fused) " gpurun_out/lt_epilogue.log; tail -1 gpurun_out/bench_ours_n1.log | cut -c1-330
