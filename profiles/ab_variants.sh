#!/bin/bash
# A/B timing of tuning builds (cron_operator_b200/csrc/variants/lib_*.so); run under gpurun.
#   lib_e*/lib_h*: unmet kernels on the C4 workload;  lib_n*: next1 kernel (C2 shape);  lib_o*: next_n (C3 shape)
V=cron_operator_b200/csrc/variants
for v in "" $(ls $V/lib_[eh]*.so 2>/dev/null); do
  echo "== unmet ${v:-default}"
  B200CRON_LIB=$v python profiles/run_kernel.py unmet 10000000 5
done
for v in "" $(ls $V/lib_n*.so 2>/dev/null); do
  echo "== next1 ${v:-default}"
  B200CRON_LIB=$v python profiles/run_kernel.py next1 1000000 200
done
for v in "" $(ls $V/lib_o*.so 2>/dev/null); do
  echo "== nextn ${v:-default}"
  B200CRON_LIB=$v python profiles/run_kernel.py nextn 8000000 5
done
if [ -n "$AB_HISTORY" ]; then
  echo "== history default"
  python profiles/run_kernel.py history 8000000 5
fi
