#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -p no:cacheprovider > gpurun_out/test_zz_alone.log 2>&1; echo "zz alone exit $?" >> gpurun_out/summary.txt
timeout 1200 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/test_all_gpu.log 2>&1; echo "all gpu tests exit $?" >> gpurun_out/summary.txt
timeout 600 python bench/optimizer_microbench.py --out gpurun_out/microbench.json > gpurun_out/microbench.log 2>&1; echo "microbench exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --steps 4 --warmup 3 > gpurun_out/bench_ours_n1.log 2>&1; echo "bench ours exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -3 gpurun_out/test_zz_alone.log; tail -3 gpurun_out/test_all_gpu.log; cat gpurun_out/microbench.log; tail -1 gpurun_out/bench_ours_n1.log | cut -c1-400
