#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 300 python bench/profile_kernels.py > gpurun_out/plain_profile.log 2>&1; echo "plain exit $?" >> gpurun_out/summary.txt
for k in lion_fused_kernel ce_kernel norm_bwd_kernel colsum_kernel norm_fwd_kernel; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 2 -c 2 -o gpurun_out/prof_$k python bench/profile_kernels.py > gpurun_out/ncu_$k.log 2>&1
  echo "ncu $k exit $?" >> gpurun_out/summary.txt
done
cat gpurun_out/summary.txt
