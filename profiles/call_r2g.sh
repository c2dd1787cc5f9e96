#!/bin/bash
# round 2, call g: full GPU suite with class A in next_n; nextn timing + launch list
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_r2g.log 2>&1
tail -8 gpurun_out/pytest_r2g.log
python profiles/run_kernel.py nextn 8000000 5 2>&1 | tee gpurun_out/ab_r2g.log
python profiles/run_kernel.py nextn 8000000 3 > gpurun_out/plain_n_r2g.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/launches_nextn_r2g.csv python profiles/run_kernel.py nextn 8000000 3 > gpurun_out/ncu_n_r2g.log 2>&1
