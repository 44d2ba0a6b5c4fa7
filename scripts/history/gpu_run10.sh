#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 300 python bench/model_kernels_microbench.py --out gpurun_out/model_kernels.json > gpurun_out/mk_default.log 2>&1; echo "default exit $?" >> gpurun_out/summary.txt
for cfg in "1 8" "1 16" "2 16" "2 4"; do set -- $cfg
  DLION_NORM_BWD_ROWS=$1 DLION_NORM_BWD_PARTS_PER_SM=$2 timeout 300 python bench/model_kernels_microbench.py 2>&1 | grep norm_bwd > gpurun_out/mk_rows$1_parts$2.log
done
timeout 600 python bench.py --steps 4 --warmup 3 --no-e2e > gpurun_out/bench_ours_n1.log 2>&1; echo "bench exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; cat gpurun_out/mk_default.log; for f in gpurun_out/mk_rows*; do echo $f; cat $f; done; tail -1 gpurun_out/bench_ours_n1.log | cut -c1-250
