#!/bin/bash
# chained-call shortcut in the UTC instantiation: whole GPU suite, then next_n A/B
( time timeout 1000 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_r2s.log 2>&1; tail -4 gpurun_out/pytest_r2s.log
{ echo "== nextn default (shortcut)"; python profiles/run_kernel.py nextn 8000000 5; echo "== nextn shortcut off"; B200CRON_LIB=cron_operator_b200/csrc/variants/lib_o_c0.so python profiles/run_kernel.py nextn 8000000 5; } 2>&1 | tee gpurun_out/ab_r2s.log
