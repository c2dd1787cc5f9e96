#!/bin/bash
# round 2, call e: GPU suite with the tier-0 zoned path; A/B against tier 0 off; full bench line
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_r2e.log 2>&1
tail -8 gpurun_out/pytest_r2e.log
bash profiles/ab_variants.sh > gpurun_out/ab_r2e.log 2>&1
cat gpurun_out/ab_r2e.log
AB_HISTORY=1 python profiles/run_kernel.py history 8000000 5
( time python bench.py ) > gpurun_out/bench_r2e.json 2> gpurun_out/bench_r2e.err
tail -c 800 gpurun_out/bench_r2e.err; head -c 5000 gpurun_out/bench_r2e.json
