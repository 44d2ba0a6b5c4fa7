#!/bin/bash
# short check of HEAD on one B200: smoke(), all GPU tests, a short bench of our arm
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/verify_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/verify_smoke.log; tail -2 gpurun_out/verify_smoke.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/verify_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/verify_pytest.log; tail -3 gpurun_out/verify_pytest.log
timeout 300 python bench.py > gpurun_out/verify_bench_ours.json 2> gpurun_out/verify_bench_ours.err; cut -c1-260 gpurun_out/verify_bench_ours.json
