#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 600 python -m pytest tests/test_multigpu.py -q -s -m gpu --maxfail=3 -p no:cacheprovider -k "ipc or auto or (majority and bfloat16 and None)" > gpurun_out/test_multigpu_ipc.log 2>&1
echo "multigpu ipc/long exit $?" >> gpurun_out/summary.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29741 bench.py --gpus 2 > gpurun_out/bench_ours_n2.log 2>&1
echo "bench n2 exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -4 gpurun_out/test_multigpu_ipc.log; tail -1 gpurun_out/bench_ours_n2.log | cut -c1-300
