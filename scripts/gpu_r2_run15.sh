#!/bin/bash
# final check of HEAD on one B200: the register-prefetch build of the MLP-backward kernel (test + microbench, all three
# builds), then the standing verify script, the evaluation-forward microbench and an ncu capture of the default build
mkdir -p gpurun_out
for v in pair staged; do
  echo "== variant $v"
  DLION_DGELU_VARIANT=$v timeout 120 python -m pytest tests/test_gpu_fused_mlp.py -x -q -m gpu -k "dgrad or node" 2>&1 | tail -2
  DLION_DGELU_VARIANT=$v timeout 60 python bench/fused_mlp_microbench.py --iters 10 --out gpurun_out/fused_mlp_mb_$v.json 2>&1 | cut -c1-260 | head -3
done
bash scripts/gpu_verify.sh 2>/dev/null | grep -v "^+"
timeout 120 python bench/eval_loss_microbench.py > gpurun_out/eval_loss_mb.json 2> gpurun_out/eval_loss_mb.err; tail -1 gpurun_out/eval_loss_mb.json; tail -2 gpurun_out/eval_loss_mb.err
timeout 300 ncu --set full --clock-control none --import-source on -k regex:dgrad_dgelu --launch-skip 2 --launch-count 1 -o gpurun_out/ncu_dgelu2 -f python bench/profile_fused_mlp.py > gpurun_out/ncu_dgelu2.log 2>&1
python scripts/ncu_summary.py gpurun_out/ncu_dgelu2.ncu-rep gpurun_out/ncu_dgrad_dgelu_kernel_r2.md "dgrad_dgelu kernel (default build), GPT-2 shape (20480 x 3072 x 768, bf16)"
