#!/bin/bash
# round 2, call l: next_n zoned launch occupancy / outlined step emulation A/B
{
for v in "" cron_operator_b200/csrc/variants/lib_o_z8.so cron_operator_b200/csrc/variants/lib_o_z12.so cron_operator_b200/csrc/variants/lib_o_ol8.so cron_operator_b200/csrc/variants/lib_o_ol12.so; do
  echo "== nextn ${v:-default}"; B200CRON_LIB=$v python profiles/run_kernel.py nextn 8000000 5
done
} 2>&1 | tee gpurun_out/ab_r2l.log
