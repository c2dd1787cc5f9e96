#!/bin/bash
# round 2, call f: unmet with class A (UTC + single-period zoned on the period clock): unmet tests, A/B of the two plans, launch list
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -q -k "unmet or C4 or smoke or reconciler or resident or every" ) > gpurun_out/pytest_r2f.log 2>&1
tail -8 gpurun_out/pytest_r2f.log
for v in "" cron_operator_b200/csrc/variants/lib_e_p0.so; do echo "== unmet ${v:-default}"; B200CRON_LIB=$v python profiles/run_kernel.py unmet 10000000 5; done 2>&1 | tee gpurun_out/ab_r2f.log
python profiles/run_kernel.py unmet 10000000 3 > gpurun_out/plain_u_r2f.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/launches_unmet_r2f.csv python profiles/run_kernel.py unmet 10000000 3 > gpurun_out/ncu_u_r2f.log 2>&1
