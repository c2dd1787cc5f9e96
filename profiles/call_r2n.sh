#!/bin/bash
# round 2, last call: smoke + a quick slice of the GPU suite at HEAD; the list emission kernel timed and captured
mkdir -p gpurun_out
python __graft_entry__.py smoke > gpurun_out/smoke_r2n.log 2>&1; tail -2 gpurun_out/smoke_r2n.log
( time timeout 600 python -m pytest tests -m gpu -q -k "golden or reference_shapes or upstream or every_k or list or zz or static" ) > gpurun_out/pytest_r2n.log 2>&1
tail -4 gpurun_out/pytest_r2n.log
L="python profiles/run_kernel.py list 2000000 3"
$L 2>&1 | tee gpurun_out/plain_list_r2n.log &&
ncu --set full --clock-control none --import-source on -k regex:unmet_list_emit -s 1 -c 1 -f -o gpurun_out/prof_list_emit_r2n $L > gpurun_out/ncu_list_emit_r2n.log 2>&1
ls -la gpurun_out/*r2n*
