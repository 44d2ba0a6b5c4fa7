#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/test_all_gpu.log 2>&1; echo "all gpu tests exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --steps 1 --warmup 1 --ga 1 --no-e2e > gpurun_out/plain_small.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches3.csv \
   python bench.py --steps 1 --warmup 1 --ga 1 --no-e2e > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?" >> gpurun_out/summary.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'lion_fused_kernel|ce_kernel|norm_bwd_kernel|norm_fwd_kernel|colsum_kernel' -c 14 \
   -o gpurun_out/prof_step python bench/profile_kernels.py > gpurun_out/ncu_step.log 2>&1
echo "ncu full exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -3 gpurun_out/test_all_gpu.log
