#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 300 python bench/profile_quant4.py > gpurun_out/plain_quant4.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:q4_dequantize_kernel -s 1 -c 1 -o gpurun_out/prof_q4_dequantize python bench/profile_quant4.py > gpurun_out/ncu_quant4.log 2>&1
echo "ncu quant4 exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; tail -2 gpurun_out/plain_quant4.log; tail -3 gpurun_out/ncu_quant4.log
