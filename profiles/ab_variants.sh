#!/bin/bash
# A/B timing of tuning builds (cron_operator_b200/csrc/variants/lib_*.so); run under gpurun.
#   lib_e*/lib_r*/lib_h*: unmet kernel pair on the C4 workload;  lib_n*: next1 kernel through bench.py (C2)
mkdir -p gpurun_out
V=cron_operator_b200/csrc/variants
for v in "" $(ls $V/lib_[erh]*.so 2>/dev/null); do
  echo "== unmet ${v:-default}"
  B200CRON_LIB=$v python profiles/run_kernel.py unmet 10000000 5
done
echo "== unmet single"
B200CRON_UNMET_SINGLE=1 python profiles/run_kernel.py unmet 10000000 5
for v in "" $(ls $V/lib_n*.so 2>/dev/null); do
  for ov in "" "--no-overlap"; do
    echo "== next1 ${v:-default} $ov"
    B200CRON_LIB=$v python bench.py --steps 100 --warmup 5 --no-extra --no-cpu $ov 2>/dev/null | grep -o '"ms_per_step": [0-9.]*, "higher\|"frac": [0-9.]*\|"e2e": {"value": [0-9.]*'
  done
done
for v in "" $(ls $V/lib_o*.so 2>/dev/null); do
  echo "== nextn ${v:-default}"
  B200CRON_LIB=$v python profiles/run_kernel.py nextn 8000000 5
done
