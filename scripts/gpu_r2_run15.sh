#!/bin/bash
# final check of HEAD on one B200: verify script, evaluation-forward microbench, ncu capture of the final MLP-backward kernel
bash scripts/gpu_verify.sh 2>/dev/null | grep -v "^+"
timeout 120 python bench/eval_loss_microbench.py > gpurun_out/eval_loss_mb.json 2> gpurun_out/eval_loss_mb.err; tail -1 gpurun_out/eval_loss_mb.json; tail -2 gpurun_out/eval_loss_mb.err
timeout 300 ncu --set full --clock-control none --import-source on -k regex:dgrad_dgelu_kernel --launch-skip 2 --launch-count 1 -o gpurun_out/ncu_dgelu2 -f python bench/profile_fused_mlp.py > gpurun_out/ncu_dgelu2.log 2>&1
python scripts/ncu_summary.py gpurun_out/ncu_dgelu2.ncu-rep gpurun_out/ncu_dgrad_dgelu_kernel_r2.md "dgrad_dgelu_kernel (final single-CTA build), GPT-2 shape (20480 x 3072 x 768, bf16)"
