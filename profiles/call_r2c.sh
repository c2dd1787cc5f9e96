#!/bin/bash
# round 2, call c: GPU suite (ring kernel, class-split next_n, spec sets, year index, @every); A/B; launch lists
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_r2c.log 2>&1
tail -25 gpurun_out/pytest_r2c.log
bash profiles/ab_variants.sh > gpurun_out/ab_r2c.log 2>&1
cat gpurun_out/ab_r2c.log
python profiles/run_kernel.py nextn 8000000 3 > gpurun_out/plain_n_r2c.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --csv --log-file gpurun_out/launches_nextn_r2c.csv python profiles/run_kernel.py nextn 8000000 3 > gpurun_out/ncu_n_r2c.log 2>&1
python profiles/run_kernel.py unmet 10000000 3 > gpurun_out/plain_u_r2c.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --csv --log-file gpurun_out/launches_unmet_r2c.csv python profiles/run_kernel.py unmet 10000000 3 > gpurun_out/ncu_u_r2c.log 2>&1
( time python bench.py --steps 5 --warmup 3 ) > gpurun_out/bench_r2c.json 2> gpurun_out/bench_r2c.err
tail -c 1500 gpurun_out/bench_r2c.err; head -c 6000 gpurun_out/bench_r2c.json
