#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/test_all_gpu.log 2>&1; echo "all gpu tests exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --steps 4 --warmup 3 > gpurun_out/bench_ours_n1.log 2>&1; echo "bench exit $?" >> gpurun_out/summary.txt
DLION_CE_STREAMING=1 timeout 600 python bench.py --steps 4 --warmup 3 --no-e2e > gpurun_out/bench_ce_streaming.log 2>&1; echo "bench ce streaming exit $?" >> gpurun_out/summary.txt
timeout 300 python bench/model_kernels_microbench.py --out gpurun_out/model_kernels.json > gpurun_out/mk_default.log 2>&1; echo "mk exit $?" >> gpurun_out/summary.txt
timeout 300 python bench/profile_kernels.py > gpurun_out/plain_profile.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lion_fused_kernel -c 2 -o gpurun_out/prof_lion_fused_kernel python bench/profile_kernels.py > gpurun_out/ncu_fused.log 2>&1
echo "ncu fused exit $?" >> gpurun_out/summary.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'ce_reg_kernel|norm_bwd_kernel' -s 4 -c 4 -o gpurun_out/prof_ce_norm python bench/profile_kernels.py > gpurun_out/ncu_ce_norm.log 2>&1
echo "ncu ce/norm exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -3 gpurun_out/test_all_gpu.log; tail -1 gpurun_out/bench_ours_n1.log | cut -c1-250; tail -1 gpurun_out/bench_ce_streaming.log | cut -c1-250; grep -E "ce_fwd|norm_bwd" gpurun_out/mk_default.log
