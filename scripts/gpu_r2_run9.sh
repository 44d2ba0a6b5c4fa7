#!/bin/bash
# round 2, call 9 (1 GPU): ncu --set full over every kernel; summaries are made on the box (the report is > 64 MiB)
set -x
mkdir -p gpurun_out
K='regex:^(ce_|clip_|colsum|embed_|gelu_|lion_|multi_gather|norm_|swiglu|rope|sumsq|sign_pack|vote_update|q4_)'
timeout 600 python bench/profile_all_kernels.py > gpurun_out/r2_9_profile_plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k "$K" -o /tmp/prof_all_r2 -f python bench/profile_all_kernels.py > gpurun_out/r2_9_ncu.log 2>&1
echo "ncu exit $?"
python scripts/ncu_summary_all.py /tmp/prof_all_r2.ncu-rep gpurun_out/ncu_all_kernels_r2.md "ncu --set full: every hand-written kernel at its training-path shape (round 2 build)"
# the fused kernel's two launches as a small report of their own + the detailed summary
ncu -i /tmp/prof_all_r2.ncu-rep --page raw --csv > /tmp/raw.csv 2>/dev/null; wc -c /tmp/raw.csv
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lion_fused_kernel -o gpurun_out/prof_lion_fused_r2 -f python bench/profile_all_kernels.py > gpurun_out/r2_9_ncu_fused.log 2>&1
python scripts/ncu_summary.py gpurun_out/prof_lion_fused_r2.ncu-rep gpurun_out/ncu_lion_fused_kernel_r2.md "ncu: lion_fused_kernel with clip + zero-fill folded in, 8 emulated ranks, 124M bf16 (round 2 build)"
ls -la gpurun_out/*.ncu-rep gpurun_out/*.md
