#!/bin/bash
# round 2, call k: GPU suite; next1 with the 5-field warp vote (six planes) vs without; next_n with the exact class-A test
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_r2k.log 2>&1
tail -6 gpurun_out/pytest_r2k.log
{
echo "== next1 six planes (warp vote)"; SIX=1 python profiles/run_kernel.py next1 1000000 200
echo "== next1 six planes, vote off"; SIX=1 B200CRON_LIB=cron_operator_b200/csrc/variants/lib_n_vote0.so python profiles/run_kernel.py next1 1000000 200
echo "== next1 five"; python profiles/run_kernel.py next1 1000000 200
echo "== nextn"; python profiles/run_kernel.py nextn 8000000 5
echo "== unmet five"; FIVE=1 python profiles/run_kernel.py unmet 10000000 5
echo "== unmet five, ordered kernel runs the general machine (no hard kernel)"; FIVE=1 B200CRON_LIB=cron_operator_b200/csrc/variants/lib_e_os1.so python profiles/run_kernel.py unmet 10000000 5
echo "== same, 6 CTAs/SM"; FIVE=1 B200CRON_LIB=cron_operator_b200/csrc/variants/lib_e_os1e6.so python profiles/run_kernel.py unmet 10000000 5
echo "== nextn zoned minblocks 8"; B200CRON_LIB=cron_operator_b200/csrc/variants/lib_o_z8.so python profiles/run_kernel.py nextn 8000000 5
} 2>&1 | tee gpurun_out/ab_r2k.log
python profiles/run_kernel.py nextn 8000000 3 > gpurun_out/plain_n_r2k.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --csv --log-file gpurun_out/launches_nextn_r2k.csv python profiles/run_kernel.py nextn 8000000 3 > gpurun_out/ncu_n_r2k.log 2>&1
