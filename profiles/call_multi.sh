#!/bin/bash
# multi-GPU evidence on N GPUs of one box: bash profiles/call_multi.sh N [tests] [ref] [bench] [pern]   (under gpurun --gpus N)
N=$1; shift
WHAT=" ${*:-tests ref bench pern} "
mkdir -p gpurun_out
nvidia-smi -L | tee gpurun_out/gpus_n$N.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
if [[ "$WHAT" == *" tests "* ]]; then
  ( time timeout 1500 python -m pytest tests/test_multi_gpu.py "tests/test_gpu_parity.py::test_multi_device_context_matches_single" -m gpu -q ) > gpurun_out/pytest_multi_n$N.log 2>&1
  tail -6 gpurun_out/pytest_multi_n$N.log
fi
if [[ "$WHAT" == *" ref "* ]]; then
  timeout 600 $TR bench.py --impl reference --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err
fi
if [[ "$WHAT" == *" bench "* ]]; then
  timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
  tail -c 400 gpurun_out/bench_n$N.err; head -c 1200 gpurun_out/bench_n$N.json
fi
if [[ "$WHAT" == *" pern "* ]]; then
  timeout 700 bash profiles/ncu_rank0.sh $N
  python profiles/pern_summary.py $N | tee gpurun_out/r2_ncu_n$N.md
fi
