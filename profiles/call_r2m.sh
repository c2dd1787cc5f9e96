#!/bin/bash
# round 2, call m: occupancy A/B of the class-A kernels (unmet tile, next_n class A)
{
for v in "" $(ls cron_operator_b200/csrc/variants/lib_e_*.so); do echo "== unmet five ${v:-default}"; FIVE=1 B200CRON_LIB=$v python profiles/run_kernel.py unmet 10000000 5; done
for v in "" $(ls cron_operator_b200/csrc/variants/lib_o_*.so); do echo "== nextn ${v:-default}"; B200CRON_LIB=$v python profiles/run_kernel.py nextn 8000000 5; done
} 2>&1 | tee gpurun_out/ab_r2m.log
