import sys, time, numpy as np
sys.path.insert(0, ".")
import torch
from cron_operator_b200 import api, synth
ctx = api.Context(devices=[0])
n = 1_000_000
m = synth.gen_specs(n); f = synth.gen_times(n, synth.ts(2024), synth.ts(2034))
hm, hf, ho = api.PinnedArray((6, n), np.uint64), api.PinnedArray(n, np.int64), api.PinnedArray((n, 1), np.int64)
hm.array[:] = m; hf.array[:] = f
ref = ctx.next_n(hm.array, hf.array, 1).copy()
L = api.lib()
planes = api._planes(hm.array)[0]
def zc():
    rc = L.b200cron_next_n_dev(ctx._h, 0, None, planes, None, hf.array.ctypes.data, n, 1, ho.array.ctypes.data)
    assert rc == 0
    ctx.sync_check()
for _ in range(5): zc()
assert np.array_equal(ho.array, ref)
t0 = time.perf_counter()
for _ in range(50): zc()
t1 = time.perf_counter()
for _ in range(5): ctx.next_n(hm.array, hf.array, 1, out=ho.array)
t2 = time.perf_counter()
for _ in range(50): ctx.next_n(hm.array, hf.array, 1, out=ho.array)
t3 = time.perf_counter()
print("zero-copy ms/step %.3f   staged ms/step %.3f" % ((t1 - t0) / 50 * 1e3, (t3 - t2) / 50 * 1e3))
