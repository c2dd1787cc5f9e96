#!/bin/bash
# per-N HBM evidence: bash profiles/ncu_rank0.sh N -- one plain process per GPU, rank 0 under ncu (three cheap metrics),
# ranks 1..N-1 keep their GPUs busy with the same work for BUSY_SECONDS while rank 0's kernels are measured
N=${1:-1}
mkdir -p gpurun_out
for r in $(seq 1 $((N - 1))); do
  BUSY_SECONDS=${BUSY_SECONDS:-50} python profiles/pern.py $r $N > gpurun_out/pern_n${N}_rank$r.log 2>&1 &
done
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
  -k regex:"next1_kernel|history_prune_kernel" --csv --log-file gpurun_out/ncu_pern_n$N.csv python profiles/pern.py 0 $N \
  > gpurun_out/pern_n${N}_rank0.log 2>&1
wait
tail -n 2 gpurun_out/pern_n${N}_rank*.log
