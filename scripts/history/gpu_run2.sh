#!/bin/bash
# Second single-GPU pass: re-verify kernels after the fused-kernel rewrite, microbench, ncu full captures.
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
for f in test_gpu_fused test_gpu_lion_kernels test_gpu_model_kernels; do
  timeout 900 python -m pytest tests/$f.py -q -s -m gpu --maxfail=40 -p no:cacheprovider > gpurun_out/$f.log 2>&1
  echo "$f exit $?" >> gpurun_out/summary.txt
done
timeout 600 python bench/optimizer_microbench.py --out gpurun_out/microbench.json > gpurun_out/microbench.log 2>&1; echo "microbench exit $?" >> gpurun_out/summary.txt
timeout 600 python bench/optimizer_microbench.py --iters 2 > gpurun_out/plain_mb.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'lion_fused_kernel|lion_local_kernel|vote_update_kernel' -c 8 \
   -o gpurun_out/prof_lion python bench/optimizer_microbench.py --iters 2 > gpurun_out/ncu_lion.log 2>&1
echo "ncu lion exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; cat gpurun_out/microbench.log
