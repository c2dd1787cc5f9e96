#!/bin/bash
# multi-GPU evidence on N GPUs of one box: bash profiles/call_multi.sh N   (run under gpurun --gpus N)
N=$1
mkdir -p gpurun_out
nvidia-smi -L | tee gpurun_out/gpus_n$N.txt
if [ "$N" = "2" ]; then
  ( time timeout 1500 python -m pytest tests/test_multi_gpu.py "tests/test_gpu_parity.py::test_multi_device_context_matches_single" -m gpu -q ) > gpurun_out/pytest_multi_n$N.log 2>&1
  tail -6 gpurun_out/pytest_multi_n$N.log
fi
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 900 $TR bench.py --impl reference --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err
timeout 1500 $TR bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
tail -c 600 gpurun_out/bench_n$N.err; head -c 2500 gpurun_out/bench_n$N.json
timeout 900 $TR --no-python profiles/ncu_rank0.sh > gpurun_out/pern_n$N.log 2>&1
tail -4 gpurun_out/pern_n$N.log
