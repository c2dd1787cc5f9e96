#!/bin/bash
# even shares for the divergence-serialised launches (unmet hard kernel, next_n zoned launch): full GPU suite + timing
( time timeout 1500 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_r2q.log 2>&1; tail -4 gpurun_out/pytest_r2q.log
{ echo "== unmet five"; FIVE=1 python profiles/run_kernel.py unmet 10000000 5; echo "== unmet 4M"; FIVE=1 python profiles/run_kernel.py unmet 4000000 5; echo "== nextn"; python profiles/run_kernel.py nextn 8000000 5; } 2>&1 | tee gpurun_out/ab_r2q.log
