#!/bin/bash
# torchrun --no-python wrapper: rank 0 under ncu (three cheap metrics: no multi-pass replay storm), the others plain
if [ "${LOCAL_RANK:-0}" = "0" ]; then
  exec ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
    -k regex:"next1_kernel|history_prune_kernel" --csv --log-file gpurun_out/ncu_pern_n${WORLD_SIZE:-1}.csv python profiles/pern.py
else
  exec python profiles/pern.py
fi
