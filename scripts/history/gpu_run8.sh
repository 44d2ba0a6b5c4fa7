#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 900 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/test_all_gpu.log 2>&1; echo "all gpu tests exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --steps 4 --warmup 3 > gpurun_out/bench_ours_n1.log 2>&1; echo "bench ours exit $?" >> gpurun_out/summary.txt
DLION_CE_CHUNK=1024 timeout 600 python bench.py --steps 4 --warmup 3 --no-e2e > gpurun_out/bench_ours_ce1024.log 2>&1; echo "bench ce1024 exit $?" >> gpurun_out/summary.txt
DLION_CE_CHUNK=2048 timeout 600 python bench.py --steps 4 --warmup 3 --no-e2e > gpurun_out/bench_ours_ce2048.log 2>&1; echo "bench ce2048 exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -4 gpurun_out/test_all_gpu.log; for f in bench_ours_n1 bench_ours_ce1024 bench_ours_ce2048; do tail -1 gpurun_out/$f.log | cut -c1-260; done
