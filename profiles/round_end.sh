#!/bin/bash
# Round-end measurement on one B200 (run under gpurun): the default bench line, then -- each after the same command
# exited 0 without ncu -- the launch list and one full capture of the headline kernel, and the same for the unmet pair.
T=${1:-r1g}
mkdir -p gpurun_out
python bench.py > gpurun_out/bench_$T.json 2> gpurun_out/bench_$T.err
SHORT="python bench.py --steps 20 --warmup 8 --no-extra --no-cpu"
$SHORT > gpurun_out/plain_$T.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$T.csv $SHORT > gpurun_out/ncu_list_$T.log 2>&1
$SHORT > gpurun_out/plain2_$T.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:next1_kernel -s 10 -c 2 -f -o gpurun_out/prof_next1_$T $SHORT > gpurun_out/ncu_full_$T.log 2>&1
if [ -n "$ROUND_END_ALL" ]; then
U="python profiles/run_kernel.py unmet 4000000"
$U > gpurun_out/plain_u_$T.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:unmet -s 6 -c 2 -f -o gpurun_out/prof_unmet_$T $U > gpurun_out/ncu_u_$T.log 2>&1
N="python profiles/run_kernel.py nextn 4000000"
$N > gpurun_out/plain_n_$T.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:next_n_kernel -s 3 -c 1 -f -o gpurun_out/prof_nextn_$T $N > gpurun_out/ncu_n_$T.log 2>&1
fi
tail -c 600 gpurun_out/bench_$T.err; head -c 1500 gpurun_out/bench_$T.json
