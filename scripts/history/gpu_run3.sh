#!/bin/bash
# 2-GPU pass: re-verify kernels, ncu the fused kernel, then the real multi-GPU path.
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/gpus.txt
nvidia-smi topo -m >> gpurun_out/gpus.txt 2>&1
export CUDA_VISIBLE_DEVICES_ALL=$CUDA_VISIBLE_DEVICES
for f in test_gpu_fused test_gpu_lion_kernels test_gpu_model_kernels; do
  timeout 600 python -m pytest tests/$f.py -q -m gpu --maxfail=20 -p no:cacheprovider > gpurun_out/$f.log 2>&1
  echo "$f exit $?" >> gpurun_out/summary.txt
done
timeout 600 python bench/optimizer_microbench.py --out gpurun_out/microbench.json > gpurun_out/microbench.log 2>&1; echo "microbench exit $?" >> gpurun_out/summary.txt
timeout 900 python -m pytest tests/test_multigpu.py -q -s -m gpu --maxfail=3 -p no:cacheprovider > gpurun_out/test_multigpu.log 2>&1
echo "test_multigpu exit $?" >> gpurun_out/summary.txt
NCCL_DEBUG=WARN timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29701 \
   bench/sign_exchange_microbench.py --sizes 1,16,124 --out gpurun_out/sign_exchange_w2.json > gpurun_out/sign_exchange_w2.log 2>&1
echo "sign_exchange exit $?" >> gpurun_out/summary.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29702 \
   bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_ours_n2.log 2>&1; echo "bench ours n2 exit $?" >> gpurun_out/summary.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29703 \
   bench.py --gpus 2 --steps 3 --warmup 3 --impl reference > gpurun_out/bench_ref_n2.log 2>&1; echo "bench ref n2 exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_ours_n1.log 2>&1; echo "bench ours n1 exit $?" >> gpurun_out/summary.txt
CUDA_VISIBLE_DEVICES=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:lion_fused_kernel -c 3 \
   -o gpurun_out/prof_fused python bench/optimizer_microbench.py --iters 1 > gpurun_out/ncu_fused.log 2>&1
echo "ncu fused exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; cat gpurun_out/microbench.log; tail -4 gpurun_out/sign_exchange_w2.log; tail -3 gpurun_out/test_multigpu.log; tail -1 gpurun_out/bench_ours_n2.log; tail -1 gpurun_out/bench_ref_n2.log
