#!/bin/bash
# round 2, call 4 (1 GPU): fused bias-gradient kernels vs separate passes, K = 1 / 2 / 4 rows in flight
set -x
mkdir -p gpurun_out
for K in 1 2 4; do DLION_COLSUM_K=$K timeout 300 python bench/bias_grad_fusion_microbench.py --out gpurun_out/r2_4_biasfusion_k$K.json 2>&1 | tail -1; done
timeout 600 python -m pytest tests/test_gpu_bias_grad_fusion.py tests/test_gpu_model_kernels.py -m gpu -x -q 2>&1 | tail -3
