#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 300 python bench/profile_kernels.py > gpurun_out/plain_profile.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:norm_fwd_kernel -s 4 -c 2 -o gpurun_out/prof_norm_fwd_v2 python bench/profile_kernels.py > gpurun_out/ncu_norm_fwd.log 2>&1
echo "ncu norm_fwd exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -2 gpurun_out/ncu_norm_fwd.log
