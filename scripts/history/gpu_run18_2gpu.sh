#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 600 python -m pytest tests/test_multigpu.py -q -m gpu --maxfail=3 -p no:cacheprovider > gpurun_out/test_multigpu2.log 2>&1
echo "multigpu W=2 exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -3 gpurun_out/test_multigpu2.log
