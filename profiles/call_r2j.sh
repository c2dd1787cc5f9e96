#!/bin/bash
# round 2, call j (1 GPU): per-N ncu at N=1; next_n launch list + ncu full of the zoned launch
bash profiles/call_multi.sh 1 pern
python profiles/run_kernel.py nextn 8000000 3 > gpurun_out/plain_n_r2j.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/launches_nextn_r2j.csv python profiles/run_kernel.py nextn 8000000 3 > gpurun_out/ncu_n_r2j.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"next_n_tile_kernel<16, false>|next_n_tile_kernel<\(int\)16, \(bool\)0>" -s 2 -c 1 -f -o gpurun_out/prof_nextn_zoned_r2j python profiles/run_kernel.py nextn 8000000 3 > gpurun_out/ncu_nz_r2j.log 2>&1
ls -la gpurun_out/*r2j*
