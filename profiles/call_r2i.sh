#!/bin/bash
# round 2, call i: full GPU suite after the zoned next_n tile change and 32-bit next1 indexing; nextn/next1 timing
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_r2i.log 2>&1
tail -6 gpurun_out/pytest_r2i.log
{
echo "== next1 five"; python profiles/run_kernel.py next1 1000000 200
echo "== next1 six planes"; SIX=1 python profiles/run_kernel.py next1 1000000 200
echo "== nextn"; python profiles/run_kernel.py nextn 8000000 5
} 2>&1 | tee gpurun_out/ab_r2i.log
