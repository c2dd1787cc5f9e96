#!/bin/bash
# round 2, call d: GPU suite; ncu --set full of the zoned kernels (unmet ordered/hard, next_n tile zoned/utc), ring variant
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_r2d.log 2>&1
tail -8 gpurun_out/pytest_r2d.log
bash profiles/ab_variants.sh > gpurun_out/ab_r2d.log 2>&1
cat gpurun_out/ab_r2d.log
U="python profiles/run_kernel.py unmet 4000000"
N="python profiles/run_kernel.py nextn 4000000"
$U > gpurun_out/plain_u_r2d.log 2>&1 && {
ncu --set full --clock-control none --import-source on -k regex:unmet_ordered -s 2 -c 1 -f -o gpurun_out/prof_unmet_ordered_r2d $U > gpurun_out/ncu_uo_r2d.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:unmet_hard -s 2 -c 1 -f -o gpurun_out/prof_unmet_hard_r2d $U > gpurun_out/ncu_uh_r2d.log 2>&1
}
$N > gpurun_out/plain_n_r2d.log 2>&1 && {
ncu --set full --clock-control none --import-source on -k regex:next_n_tile -s 4 -c 2 -f -o gpurun_out/prof_nextn_tile_r2d $N > gpurun_out/ncu_nt_r2d.log 2>&1
}
B200CRON_LIB=cron_operator_b200/csrc/variants/lib_n_ring1.so python profiles/run_kernel.py next1 1000000 20 > gpurun_out/plain_ring_r2d.log 2>&1 && \
B200CRON_LIB=cron_operator_b200/csrc/variants/lib_n_ring1.so ncu --set full --clock-control none --import-source on -k regex:next1_ring -s 40 -c 1 -f -o gpurun_out/prof_ring_r2d python profiles/run_kernel.py next1 1000000 20 > gpurun_out/ncu_ring_r2d.log 2>&1
ls -la gpurun_out/*.ncu-rep
