"""Small driver used for ncu captures of the non-headline kernels:  python profiles/run_kernel.py {unmet|nextn|history} [n]"""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from cron_operator_b200 import api, synth  # noqa: E402

import os  # noqa: E402

api.lib(os.environ.get("B200CRON_LIB") or None)  # A/B: a tuning build from profiles/build_variants.sh
which = sys.argv[1]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2_000_000
ctx = api.Context(devices=[0])
names = ctx.zone_names
dst = np.array([z for z in range(1, len(names)) if (ctx.zone_periods(z)[0] > synth.ts(2030)).any()])
stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)
st = stream.cuda_stream
t = lambda a: torch.from_numpy(a).cuda()  # noqa: E731
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
if which == "next1":
    # the headline shape without bench.py's host-side set-up: 8 device-generated 1M-spec UTC batches, K back-to-back
    # launches, overlap on and off (argv[3] = K)
    K = int(sys.argv[3]) if len(sys.argv) > 3 else 200
    nb, n1 = 8, 1_000_000
    dm = [torch.empty((6, n1), dtype=torch.int64, device="cuda") for _ in range(nb)]
    df = [torch.empty(n1, dtype=torch.int64, device="cuda") for _ in range(nb)]
    do = [torch.empty((n1, 1), dtype=torch.int64, device="cuda") for _ in range(nb)]
    for b in range(nb):
        ctx.synth_specs_dev(dm[b], n1, 0xC0FFEE, b * n1, False, stream=st)
        ctx.synth_times_dev(df[b], n1, synth.ts(2024), synth.ts(2034), 0xC0FFEE, b * n1, stream=st)
    planes = [api._planes(x, five_field=not os.environ.get("SIX"))[0] for x in dm]  # SIX=1: pass the seconds plane too
    L = api.lib()
    for ov in (1, 0):
        ctx.set_overlap(bool(ov))
        best = 1e30
        for rep in range(4):
            for i in range(16):
                L.b200cron_next_n_dev(ctx._h, 0, st, planes[i % nb], None, df[i % nb].data_ptr(), n1, 1, do[i % nb].data_ptr())
            ctx.sync_check(stream=st)
            ev0.record()
            for i in range(K):
                L.b200cron_next_n_dev(ctx._h, 0, st, planes[i % nb], None, df[i % nb].data_ptr(), n1, 1, do[i % nb].data_ptr())
            ev1.record()
            ctx.sync_check(stream=st)
            best = min(best, ev0.elapsed_time(ev1) / K)
        print("next1 overlap", ov, "us/step", best * 1e3, "roofline frac", 64e6 / (best * 1e-3) / 1e9 / 6536.7)
    ctx.close()
    sys.exit(0)
if which == "unmet":
    m = synth.gen_specs(n, seed=41, dense_minutes=0.5)
    z = synth.gen_zone_ids(n, len(names), dst, seed=41)
    now = synth.ts(2026, 3, 29, 12)
    e = now - 30 * 86400 + synth.gen_times(n, -86400, 86400, seed=41)
    rng = np.random.default_rng(41)
    dm, dz, de = t(m.view(np.int64)), t(z), t(e)
    dp, da = t(rng.integers(0, 3, n).astype(np.uint8)), t(rng.choice([0, 1, 2], size=n, p=[.6, .3, .1]).astype(np.uint32))
    dfl, ddl = t((rng.random(n) < 0.02).astype(np.uint8)), t(np.where(rng.random(n) < 0.01, now - 100, -2**63).astype(np.int64))
    o = dict(last=torch.empty(n, dtype=torch.int64, device="cuda"), next=torch.empty(n, dtype=torch.int64, device="cuda"),
             missed=torch.empty(n, dtype=torch.int32, device="cuda"), status=torch.empty(n, dtype=torch.uint8, device="cuda"),
             decision=torch.empty(n, dtype=torch.uint8, device="cuda"))
    five = bool(os.environ.get("FIVE"))  # FIVE=1: no seconds plane (5-field schedules)
    fn = lambda: ctx.unmet_since_dev(dm, de, now, o, 0, dp, da, dfl, ddl, dz, stream=st, five_field=five)  # noqa: E731
elif which == "list":
    # the capped list of missed runs (cap 100) on the C4 shape: offsets, then the warp-cooperative emission
    m = synth.gen_specs(n, seed=41, dense_minutes=0.5)
    z = synth.gen_zone_ids(n, len(names), dst, seed=41)
    now = synth.ts(2026, 3, 29, 12)
    e = now - 30 * 86400 + synth.gen_times(n, -86400, 86400, seed=41)
    dm, dz, de = t(m.view(np.int64)), t(z), t(e)
    o = dict(last=torch.empty(n, dtype=torch.int64, device="cuda"), next=torch.empty(n, dtype=torch.int64, device="cuda"),
             missed=torch.empty(n, dtype=torch.int32, device="cuda"), status=torch.empty(n, dtype=torch.uint8, device="cuda"),
             decision=torch.empty(n, dtype=torch.uint8, device="cuda"))
    ctx.unmet_since_dev(dm, de, now, o, zone_id=dz, stream=st)
    off = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    ctx.unmet_list_offsets_dev(o["missed"], o["decision"], off, 100, stream=st)
    ctx.sync_check(stream=st)
    total = int(off[-1].item())
    lst = torch.empty(total, dtype=torch.int64, device="cuda")
    print("list entries", total, "bytes", total * 8)
    fn = lambda: ctx.unmet_list_emit_dev(dm, de, off, lst, dz, stream=st)  # noqa: E731
elif which == "nextn":
    m = synth.gen_specs(n, seed=3, with_seconds=True)
    z = synth.gen_zone_ids(n, len(names), dst, seed=3)
    f = synth.gen_times(n, synth.ts(2024), synth.ts(2036), seed=3)
    dm, dz, df = t(m.view(np.int64)), t(z), t(f)
    o = torch.empty((n, 16), dtype=torch.int64, device="cuda")
    fn = lambda: ctx.next_n_dev(dm, df, o, 16, dz, stream=st)  # noqa: E731
else:
    rng = np.random.default_rng(9)
    lens = rng.integers(0, 17, n)
    seg = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    key = (1_760_000_000 - rng.integers(0, 86400, int(seg[-1]))).astype(np.int64)
    ds, dk, dl = t(seg.view(np.int64)), t(key), t(np.full(n, 10, np.int64))
    perm, keep = torch.empty(len(key), dtype=torch.int32, device="cuda"), torch.empty(len(key), dtype=torch.uint8, device="cuda")
    fn = lambda: ctx.history_prune_dev(ds, dk, dl, perm, keep, stream=st)  # noqa: E731
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
for _ in range(3 if reps == 1 else 20):  # reps > 1: A/B timing mode, warm the clocks up first
    fn()
ctx.sync_check(stream=st)
best = 1e30
for _ in range(reps):
    ev0.record()
    fn()
    ev1.record()
    ctx.sync_check(stream=st)
    best = min(best, ev0.elapsed_time(ev1))
extra = ""
if which == "nextn":  # variants must agree: a cheap fingerprint of the whole output
    extra = " checksum %d" % int((o ^ (o >> 17)).sum().item())
if which == "list":
    extra = " checksum %d  GB/s out %.0f" % (int((lst ^ (lst >> 17)).sum().item()), total * 8 / (best * 1e-3) / 1e9)
if which == "unmet":
    extra = " checksum %d" % int(sum((v.to(torch.int64) * (j + 3)).sum().item() for j, v in enumerate(o.values())) & (2**63 - 1))
print(which, n, "ms", best, extra)
ctx.close()
