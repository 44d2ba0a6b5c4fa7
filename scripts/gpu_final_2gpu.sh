#!/bin/bash
# Final 2-GPU regression: multi-GPU tests, then the headline bench on 2 GPUs (ours).
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 600 python -m pytest tests/test_multigpu.py -q -m gpu --maxfail=3 -p no:cacheprovider > gpurun_out/test_multigpu2.log 2>&1
echo "multigpu W=2 exit $?" >> gpurun_out/summary.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/bench_ours_n2.log 2>&1
echo "bench n2 exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -3 gpurun_out/test_multigpu2.log; tail -1 gpurun_out/bench_ours_n2.log | cut -c1-400
