#!/bin/bash
# round 2, call h: GPU suite; next1 five-field A/B; nextn with marked rows; unmet five-field A/B
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_r2h.log 2>&1
tail -6 gpurun_out/pytest_r2h.log
{
echo "== next1 default (5-field instantiation)"; python profiles/run_kernel.py next1 1000000 200
echo "== next1 general instantiation on the same NULL-plane launches"; B200CRON_LIB=cron_operator_b200/csrc/variants/lib_n_five0.so python profiles/run_kernel.py next1 1000000 200
echo "== next1 six planes"; SIX=1 python profiles/run_kernel.py next1 1000000 200
echo "== nextn"; python profiles/run_kernel.py nextn 8000000 5
echo "== unmet six planes"; python profiles/run_kernel.py unmet 10000000 5
echo "== unmet five planes"; FIVE=1 python profiles/run_kernel.py unmet 10000000 5
} 2>&1 | tee gpurun_out/ab_r2h.log
python profiles/run_kernel.py nextn 8000000 3 > gpurun_out/plain_n_r2h.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/launches_nextn_r2h.csv python profiles/run_kernel.py nextn 8000000 3 > gpurun_out/ncu_n_r2h.log 2>&1
