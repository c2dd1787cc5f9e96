#!/bin/bash
# resident-set call on pinned buffers: in place over PCIe (default) vs staged DMA copies
{ echo "== default (in place)"; python profiles/set_e2e.py; echo "== hybrid (DMA in, in place out)"; B200CRON_LIB=cron_operator_b200/csrc/variants/lib_s_hybrid.so python profiles/set_e2e.py; } 2>&1 | tee gpurun_out/ab_r2o.log
