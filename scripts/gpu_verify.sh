#!/bin/bash
# standing check of HEAD on one B200: build stamp, smoke(), all GPU tests, a short bench of both arms
set -x
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/verify_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/verify_smoke.log; tail -2 gpurun_out/verify_smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/verify_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/verify_pytest.log; tail -4 gpurun_out/verify_pytest.log
timeout 600 python bench.py > gpurun_out/verify_bench_ours.json 2> gpurun_out/verify_bench_ours.err; cut -c1-260 gpurun_out/verify_bench_ours.json
timeout 600 python bench.py --impl reference > gpurun_out/verify_bench_ref.json 2> gpurun_out/verify_bench_ref.err; cut -c1-260 gpurun_out/verify_bench_ref.json; tail -2 gpurun_out/verify_bench_ref.err
