#!/bin/bash
# round 2, call 10 (1 GPU): TMA norm forward + averaged-vote table — tests, microbench A/B, bench
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_10_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_10_pytest.log
tail -6 gpurun_out/r2_10_pytest.log
timeout 300 python bench/model_kernels_microbench.py --out gpurun_out/r2_10_model_kernels_tma.json > gpurun_out/r2_10_mk_tma.log 2>&1; grep -i "norm_fwd" gpurun_out/r2_10_mk_tma.log
DLION_NORM_TMA=0 timeout 300 python bench/model_kernels_microbench.py --out gpurun_out/r2_10_model_kernels_reg.json > gpurun_out/r2_10_mk_reg.log 2>&1; grep -i "norm_fwd" gpurun_out/r2_10_mk_reg.log
timeout 600 python bench.py --gpus 1 --steps 6 --warmup 3 > gpurun_out/r2_10_bench_n1.json 2> gpurun_out/r2_10_bench_n1.err; cut -c1-330 gpurun_out/r2_10_bench_n1.json; tail -3 gpurun_out/r2_10_bench_n1.err
