#!/bin/bash
# Round-end measurement on one B200 (under gpurun): GPU suite, the default bench line, then -- each after the same command
# exited 0 without ncu -- the launch list of a short bench run and one full capture of the headline kernel.
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_r2_final.log 2>&1
tail -4 gpurun_out/pytest_r2_final.log
( time python bench.py ) > gpurun_out/bench_r2_final.json 2> gpurun_out/bench_r2_final.err
tail -c 300 gpurun_out/bench_r2_final.err; head -c 600 gpurun_out/bench_r2_final.json; echo
SHORT="python bench.py --steps 1 --warmup 3 --no-extra --no-cpu"
$SHORT > gpurun_out/plain_r2_final.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r2_final.csv $SHORT > gpurun_out/ncu_list_r2_final.log 2>&1
$SHORT > gpurun_out/plain2_r2_final.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:next1_kernel -s 40 -c 2 -f -o gpurun_out/prof_next1_r2_final $SHORT > gpurun_out/ncu_full_r2_final.log 2>&1
U="python profiles/run_kernel.py unmet 4000000"
FIVE=1 $U > gpurun_out/plain_u_r2_final.log 2>&1 &&
FIVE=1 ncu --set full --clock-control none --import-source on -k regex:unmet_tile -s 2 -c 1 -f -o gpurun_out/prof_unmet_tile_r2_final $U > gpurun_out/ncu_ut_r2_final.log 2>&1
ls -la gpurun_out/*r2_final*
{
for v in "" $(ls cron_operator_b200/csrc/variants/lib_o_*.so 2>/dev/null); do echo "== nextn ${v:-default}"; B200CRON_LIB=$v python profiles/run_kernel.py nextn 8000000 5; done
} 2>&1 | tee gpurun_out/ab_r2_final.log
