#!/bin/bash
# round 2, call 2 (2 GPUs): all GPU tests incl. multi-GPU at W=2, ATen kernel attribution, bench N=1/2, microbench W=2
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_2_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_2_pytest.log
tail -8 gpurun_out/r2_2_pytest.log
timeout 300 python scripts/attribute_aten_kernels.py gpurun_out/r2_2_aten_sites.json > gpurun_out/r2_2_aten_sites.txt 2>&1; head -60 gpurun_out/r2_2_aten_sites.txt
timeout 600 python bench.py --gpus 1 --steps 6 --warmup 3 > gpurun_out/r2_2_bench_n1.json 2> gpurun_out/r2_2_bench_n1.err; cat gpurun_out/r2_2_bench_n1.json; tail -3 gpurun_out/r2_2_bench_n1.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/r2_2_bench_n2.json 2> gpurun_out/r2_2_bench_n2.err; cat gpurun_out/r2_2_bench_n2.json; tail -5 gpurun_out/r2_2_bench_n2.err
for B in 1 10; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench/sign_exchange_microbench.py --sizes 1,16,124,1000 --burst $B --out gpurun_out/r2_2_signx_w2_burst$B.json > gpurun_out/r2_2_signx_w2_burst$B.log 2>&1; tail -4 gpurun_out/r2_2_signx_w2_burst$B.log
done
