#!/bin/bash
# round 2, call b: GPU suite on the restructured unmet path + @every kind; unmet plan A/B; smoke launch list
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_r2b.log 2>&1
tail -15 gpurun_out/pytest_r2b.log
bash profiles/ab_variants.sh > gpurun_out/ab_r2b.log 2>&1
cat gpurun_out/ab_r2b.log
python __graft_entry__.py smoke > gpurun_out/smoke_r2b.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_smoke_r2b.csv python __graft_entry__.py smoke > gpurun_out/ncu_smoke_r2b.log 2>&1
tail -2 gpurun_out/smoke_r2b.log
python profiles/run_kernel.py unmet 10000000 3 > gpurun_out/plain_u_r2b.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_unmet_r2b.csv python profiles/run_kernel.py unmet 10000000 3 > gpurun_out/ncu_u_r2b.log 2>&1
