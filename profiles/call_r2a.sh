#!/bin/bash
mkdir -p gpurun_out
( which go; go version ) > gpurun_out/go_probe.txt 2>&1
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
nproc >> gpurun_out/topo.txt; lscpu | head -20 >> gpurun_out/topo.txt
bash profiles/ab_variants.sh > gpurun_out/ab_r2a.log 2>&1
python profiles/run_kernel.py unmet 10000000 3 > gpurun_out/plain_u_r2a.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_unmet_r2a.csv python profiles/run_kernel.py unmet 10000000 3 > gpurun_out/ncu_u_r2a.log 2>&1
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_r2a.log 2>&1
tail -5 gpurun_out/pytest_r2a.log; cat gpurun_out/ab_r2a.log
