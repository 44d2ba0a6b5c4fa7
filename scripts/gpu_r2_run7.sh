#!/bin/bash
# round 2, call 7 (1 GPU): tests, CE A/B, bench, ncu --set full over every kernel
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_7_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_7_pytest.log
tail -6 gpurun_out/r2_7_pytest.log
timeout 300 python bench/model_kernels_microbench.py --out gpurun_out/r2_7_model_kernels_rowres.json > gpurun_out/r2_7_mk_rowres.log 2>&1; tail -3 gpurun_out/r2_7_mk_rowres.log
DLION_CE_STREAMING=1 timeout 300 python bench/model_kernels_microbench.py --out gpurun_out/r2_7_model_kernels_streaming.json > gpurun_out/r2_7_mk_streaming.log 2>&1; tail -3 gpurun_out/r2_7_mk_streaming.log
timeout 600 python bench.py --gpus 1 --steps 6 --warmup 3 > gpurun_out/r2_7_bench_n1.json 2> gpurun_out/r2_7_bench_n1.err; cat gpurun_out/r2_7_bench_n1.json; tail -3 gpurun_out/r2_7_bench_n1.err
timeout 600 python bench/profile_all_kernels.py > gpurun_out/r2_7_profile_plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:dlion -o gpurun_out/prof_all_r2 -f python bench/profile_all_kernels.py > gpurun_out/r2_7_ncu.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/r2_7_profile_plain.log; tail -3 gpurun_out/r2_7_ncu.log; ls -la gpurun_out/prof_all_r2.ncu-rep
