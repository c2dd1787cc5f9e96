#!/bin/bash
# history_prune with the packed key+position rank loop: parity tests, then A/B against the unpacked loop
( timeout 600 python -m pytest tests -m gpu -q -k "history or golden or reference_shapes or reconciler or smoke" ) 2>&1 | tail -3
{ echo "== history default (packed)"; python profiles/run_kernel.py history 8000000 5; echo "== history unpacked"; B200CRON_LIB=cron_operator_b200/csrc/variants/lib_h_pack0.so python profiles/run_kernel.py history 8000000 5; } 2>&1 | tee gpurun_out/ab_r2p.log
