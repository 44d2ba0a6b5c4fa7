#!/bin/bash
# round 2, call 8 (1 GPU): CE rowres with 1024 threads, ncu --set full over every kernel
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_model_kernels.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python bench/model_kernels_microbench.py --out gpurun_out/r2_8_model_kernels.json > gpurun_out/r2_8_mk.log 2>&1; grep ce_fwd gpurun_out/r2_8_mk.log
timeout 600 python bench/profile_all_kernels.py > gpurun_out/r2_8_profile_plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k 'regex:^(ce_|clip_|colsum|embed_|gelu_|lion_|multi_gather|norm_|swiglu|rope|sumsq|sign_pack|vote_update|q4_)' -o gpurun_out/prof_all_r2 -f python bench/profile_all_kernels.py > gpurun_out/r2_8_ncu.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/r2_8_profile_plain.log; tail -3 gpurun_out/r2_8_ncu.log; ls -la gpurun_out/prof_all_r2.ncu-rep
