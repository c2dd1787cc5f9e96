"""Per-N HBM evidence (north_star: "ncu must show achieved HBM GB/s ... at 1, 2, 4 and 8 GPUs"): every rank runs its C5
shard -- 1e9/N UTC specs, next fire (next1_kernel<true>), and 2e7 history segments (history_prune_kernel) -- at the same
time; rank 0 runs under `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum` (see
profiles/ncu_rank0.sh), the other ranks plain and in a loop for BUSY_SECONDS so that they are busy while rank 0's
kernels are measured.  Plain processes, one per GPU, no torch.distributed: there is no collective on this path (and the
first attempt, with an NCCL barrier, hung under ncu).  usage: pern.py RANK WORLD"""
import os
import sys

import numpy as np
import torch
import time

sys.path.insert(0, ".")
from cron_operator_b200 import api, synth  # noqa: E402

rank, world = int(sys.argv[1]), int(sys.argv[2])
busy = float(os.environ.get("BUSY_SECONDS", "0"))
torch.cuda.set_device(rank)
ctx = api.Context(devices=[rank])
stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)
st = stream.cuda_stream
N5 = 1_000_000_000
lo, hi = N5 * rank // world, N5 * (rank + 1) // world
n = hi - lo
dm = torch.empty((6, n), dtype=torch.int64, device="cuda")
df = torch.empty(n, dtype=torch.int64, device="cuda")
ctx.synth_specs_dev(dm, n, 0xC0FFEE, lo, False, stream=st)
ctx.synth_times_dev(df, n, synth.ts(2024), synth.ts(2034), 0xC0FFEE, lo, stream=st)
o = torch.empty((n, 1), dtype=torch.int64, device="cuda")
ns = 20_000_000
lens = torch.empty(ns, dtype=torch.int64, device="cuda")
ctx.synth_times_dev(lens, ns, 0, 17, 0x515, lo, stream_id=301, stream=st)
ctx.sync_check(stream=st)
seg = torch.zeros(ns + 1, dtype=torch.int64, device="cuda")
torch.cumsum(lens, 0, out=seg[1:])
m = int(seg[-1].item())
key = torch.empty(m, dtype=torch.int64, device="cuda")
ctx.synth_times_dev(key, m, 1_760_000_000 - 86399, 1_760_000_001, 0x515 + rank, 0, stream_id=302, stream=st)
lim = torch.full((ns,), 10, dtype=torch.int64, device="cuda")
perm = torch.empty(m, dtype=torch.int32, device="cuda")
keep = torch.empty(m, dtype=torch.uint8, device="cuda")
ctx.sync_check(stream=st)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
t_end = time.time() + busy
rep = -1
while True:
    rep += 1
    if rep >= 3 and time.time() >= t_end:
        break
    ev[0].record()
    ctx.next_n_dev(dm, df, o, 1, None, stream=st, five_field=True)
    ev[1].record()
    ev[2].record()
    ctx.history_prune_dev(seg, key, lim, perm, keep, stream=st)
    ev[3].record()
    ctx.sync_check(stream=st)
    if rep < 3:
        print(f"rank {rank}/{world} rep {rep}: next1 {n} specs {ev[0].elapsed_time(ev[1]):.3f} ms; "
              f"history {ns} segments {m} records {ev[2].elapsed_time(ev[3]):.3f} ms", flush=True)
print(f"rank {rank}/{world}: {rep} reps", flush=True)
ctx.close()
