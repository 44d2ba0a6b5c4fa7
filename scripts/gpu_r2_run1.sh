#!/bin/bash
# round 2, call 1 (1 GPU): GPU tests, CE GEMM shape A/B, headline bench
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2_1_smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_1_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_1_pytest.log
tail -5 gpurun_out/r2_1_pytest.log
timeout 300 python scripts/ab_ce_gemm_shapes.py > gpurun_out/r2_1_ab_ce.json 2> gpurun_out/r2_1_ab_ce.err; cat gpurun_out/r2_1_ab_ce.json
timeout 600 python bench.py --gpus 1 --steps 6 --warmup 3 > gpurun_out/r2_1_bench_n1.json 2> gpurun_out/r2_1_bench_n1.err; cat gpurun_out/r2_1_bench_n1.json; tail -3 gpurun_out/r2_1_bench_n1.err
