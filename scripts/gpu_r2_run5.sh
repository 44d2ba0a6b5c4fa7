#!/bin/bash
# round 2, call 5 (1 GPU): tests, bench, launch list
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_5_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_5_pytest.log
tail -6 gpurun_out/r2_5_pytest.log
timeout 600 python bench.py --gpus 1 --steps 6 --warmup 3 > gpurun_out/r2_5_bench_n1.json 2> gpurun_out/r2_5_bench_n1.err; cat gpurun_out/r2_5_bench_n1.json; tail -3 gpurun_out/r2_5_bench_n1.err
timeout 600 python bench.py --steps 1 --warmup 1 --ga 1 --no-e2e > gpurun_out/r2_5_plain_small.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_gpt2_microstep_r2_5.csv \
   python bench.py --steps 1 --warmup 1 --ga 1 --no-e2e > gpurun_out/r2_5_ncu_launches.log 2>&1
echo "ncu exit $?"
