"""A/B of the resident-set call on pinned buffers (B200CRON_LIB = a tuning build): wall clock per b200cron_next_n_set call"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from cron_operator_b200 import api, synth  # noqa: E402

api.lib(os.environ.get("B200CRON_LIB") or None)
ctx = api.Context(devices=[0])
n = 1_000_000
masks = synth.gen_specs(n)
frm = synth.gen_times(n, synth.ts(2024), synth.ts(2034))
hf, ho = api.PinnedArray(n, np.int64), api.PinnedArray((n, 1), np.int64)
hf.array[:] = frm
st = ctx.specset(masks, None, five_field=True)
want = ctx.next_n(masks, frm, 1)
for _ in range(10):
    st.next_n(hf.array, 1, out=ho.array)
best = 1e9
for rep in range(5):
    t0 = time.perf_counter()
    for _ in range(50):
        st.next_n(hf.array, 1, out=ho.array)
    best = min(best, (time.perf_counter() - t0) / 50)
assert np.array_equal(ho.array, want), "set call disagrees with next_n"
print("set.next_n pinned: %.1f us/call, %.3e evals/s, checksum %d" % (best * 1e6, n / best, int(ho.array.sum())))
st.close()
ctx.close()
