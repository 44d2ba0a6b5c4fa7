#!/bin/bash
# Launch list (per-kernel durations) of Llama-2-7B full-parameter SFT steps: plain run first, then ncu.
mkdir -p gpurun_out; rm -f gpurun_out/summary.txt
CMD="python bench.py --model llama-2-7b --batch 4 --ga 2 --steps 1 --warmup 3 --no-e2e"
timeout 600 $CMD > gpurun_out/bench_llama7b_plain.log 2>&1; rc=$?; echo "plain exit $rc" >> gpurun_out/summary.txt
if [ $rc -eq 0 ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 4000 -c 2600 --csv --log-file gpurun_out/launches_llama7b.csv $CMD > gpurun_out/ncu_llama.log 2>&1
  echo "ncu exit $?" >> gpurun_out/summary.txt
fi
cat gpurun_out/summary.txt; tail -1 gpurun_out/bench_llama7b_plain.log | cut -c1-300; wc -l gpurun_out/launches_llama7b.csv
