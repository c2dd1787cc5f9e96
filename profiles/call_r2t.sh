#!/bin/bash
# chained-call shortcut: unmet A/B (C4 shape)
{ echo "== unmet default (shortcut)"; FIVE=1 python profiles/run_kernel.py unmet 10000000 5; echo "== unmet shortcut off"; FIVE=1 B200CRON_LIB=cron_operator_b200/csrc/variants/lib_o_c0.so python profiles/run_kernel.py unmet 10000000 5; } 2>&1 | tee gpurun_out/ab_r2t.log
