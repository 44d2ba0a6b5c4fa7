#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
timeout 600 python bench/attention_backends_microbench.py --out gpurun_out/attention_backends.json > gpurun_out/attention_backends.log 2>&1; echo "attn exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; grep -v Warning gpurun_out/attention_backends.log | tail -20
