#!/bin/bash
# Tuning builds for profiles/ab_variants.sh: name=flags pairs, e.g.
#   bash profiles/build_variants.sh e8r4="-DB200CRON_UNMET_EASY_ROUNDS=4" o_s0="-DB200CRON_NEXTN_STAGE=0"
# Macros (kernels.cuh): B200CRON_NEXT1_THREADS / _MINBLOCKS, B200CRON_NEXTN_STAGE / _MINBLOCKS,
# B200CRON_UNMET_EASY_ROUNDS / _EASY_MINBLOCKS / _HARD_MINBLOCKS / _PLAN (b200cron.cu).  Name prefixes pick the workload: e*/h* unmet (C4),
# n* next1 (C2), o* next_n (C3).  Run here (nvcc cross-compiles); the .so files travel to the GPU box with gpurun.
cd "$(dirname "$0")/../cron_operator_b200/csrc" && mkdir -p variants
for kv in "$@"; do
  name=${kv%%=*}; flags=${kv#*=}
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared $flags \
    -o variants/lib_$name.so b200cron.cu parse.cpp zone_table.cpp > /dev/null 2>&1 && echo "built variants/lib_$name.so ($flags)" &
done
wait
