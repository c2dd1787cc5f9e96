#!/usr/bin/env python
"""bench.py -- schedule evaluations/sec of the batched cron-schedule path on B200 (BASELINE.json metric).

Workload at every N: BASELINE config C2 (1M random 5-field UTC cron specs, next single fire time) per GPU -- weak
scaling: each rank owns its own contiguous shard of one big synthetic spec list; no collective on the data path.
A "step" is one pass of the hot path over STEP_BATCHES C2 batches of 1M specs each (one launch per batch, rotating
through 8 distinct batches = 512 MB, larger than L2): with the driver's K = 20 that is 3,200 launches and ~45 ms of device
work per rank between the two events, so the number measures the GPU and not launch jitter.

  value        evals/s with inputs resident in HBM: K steps between one CUDA-event pair on the launch stream, max over ranks
  e2e          the same metric through the reference-facing C ABI with pinned HOST buffers (H2D + kernel + D2H inside);
               e2e.resident = the same through a resident spec set (immutable masks stay in HBM, only times move)
  roofline     algorithmic bytes (64 B/eval) / average launch duration vs MEASURED_PEAKS.json hbm_gbs
  cpu_baseline the C restatement of the reference's Go path (oracle/) on this box's host cores (rank 0, every N)
  extra        the other BASELINE configs: C3 (100M x 16 zoned), C4 (10M Crons unmet, device + e2e + CPU sample),
               C5 sharded over the N ranks (1e9/N specs + history per GPU) with sampled oracle parity

--impl reference times that CPU path alone (the reference is Go; no Go toolchain exists on this image or on the GPU
box -- `go version`: not found -- so the reference arm is the oracle port, kind "port").
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from cron_operator_b200 import synth  # noqa: E402

N_SPECS = 1_000_000          # C2
N_BATCH = 8                  # distinct C2 batches per rank (512 MB of inputs + outputs: larger than the 126 MB L2)
STEP_BATCHES = 160           # launches per step
BYTES_PER_EVAL = 64          # 48 masks + 8 from + 8 out
T_LO, T_HI = synth.ts(2024), synth.ts(2034)
METRIC = "schedule_evals_per_sec"
UNIT = "evals/s"
WORKLOAD = "C2: 1M random 5-field UTC cron specs, next single fire time (per GPU)"
SM_ISSUE_SLOTS = 148 * 4     # warp-instruction issue slots per clock on B200
# per-launch figures taken from ncu captures of this round's kernels (profiles/r2_*): dram bytes, executed warp instr
NCU = json.load(open(os.path.join(ROOT, "profiles", "ncu_constants.json")))


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 100 ms while the timed region runs"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.rows, self.p = gpu_index, [], None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.p = None

    def _pump(self):
        for line in self.p.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.p:
            time.sleep(0.25)
            self.p.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) < 9:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference(steps: int, warmup: int, threads: int | None = None):
    """times the oracle (C restatement of the Go path) on the C2 workload with all host threads"""
    from oracle import pyoracle as po
    threads = threads or po.cpu_count()
    masks = synth.gen_specs(N_SPECS)
    frm = synth.gen_times(N_SPECS, T_LO, T_HI)
    for _ in range(warmup):
        po.batch_next_n(masks, frm, 1, threads=threads)
    t0 = time.perf_counter()
    for _ in range(steps):
        po.batch_next_n(masks, frm, 1, threads=threads)
    dt = time.perf_counter() - t0
    return N_SPECS * steps / dt, dt / steps * 1e3, threads


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    v, ms, threads = cpu_reference(args.steps, args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "note": "CPU path; every step is one full 1M-spec C2 batch on host cores (a "
                   "bounded sample of the GPU arm's step, which is 160 such batches)"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"full C2 batch ({N_SPECS} specs) x {args.steps} steps; C restatement of "
                                   "robfig/cron v3.0.1 + Go time (no Go toolchain on this image or the GPU box)"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def timed(torch, fn, reps=5, flush=None):
    """best-of-reps device time of fn() in seconds, L2 flushed before each (CUDA events on the current stream)"""
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e30
    for _ in range(reps):
        if flush is not None:
            flush.zero_()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best * 1e-3


def issue_frac(inst_per_unit, units, seconds, clocks):
    mhz = (clocks or {}).get("sm_mhz") or 1965.0
    return inst_per_unit * units / seconds / (SM_ISSUE_SLOTS * mhz * 1e6)


def c3_c4(ctx, torch, st, peak, flush, clocks):
    """kernel-only numbers for C3 and C4 at full size + C4 end to end, each with its roofline and issue fractions"""
    from cron_operator_b200 import api
    out = {}
    names = ctx.zone_names
    dst = np.array([z for z in range(1, len(names)) if (ctx.zone_periods(z)[0] > synth.ts(2030)).any()])
    # C3 at full size: 100M specs with a seconds field, mixed zones, K = 16 -- generated on the device
    n, k = 100_000_000, 16
    pool = torch.from_numpy(synth.zone_pool(len(names), dst)).cuda()
    dm = torch.empty((6, n), dtype=torch.int64, device="cuda")
    df = torch.empty(n, dtype=torch.int64, device="cuda")
    dz = torch.empty(n, dtype=torch.int32, device="cuda")
    ctx.synth_specs_dev(dm, n, 3, 0, True, stream=st)
    ctx.synth_times_dev(df, n, synth.ts(2024), synth.ts(2036), 3, 0, stream=st)
    ctx.synth_zones_dev(dz, n, pool, 3, 0, stream=st)
    o = torch.empty((n, k), dtype=torch.int64, device="cuda")
    s = timed(torch, lambda: ctx.next_n_dev(dm, df, o, k, dz, stream=st), reps=3, flush=flush)
    ctx.sync_check(stream=st)
    ipe = NCU.get("nextn16_warp_inst_per_eval")
    out["C3_nextN_k16_zoned_100M"] = {
        "specs": n, "evals_per_s": n * k / s, "specs_per_s": n / s, "ms": s * 1e3, "bytes_per_spec": 188,
        "roofline_frac": 188 * n / s / 1e9 / peak,
        "issue_frac": issue_frac(ipe, n * k, s, clocks) if ipe else None, "warp_inst_per_eval": ipe,
        "issue_source": NCU.get("nextn16_source"), "data": "generated on device (b200cron_synth_*_dev), seed 3"}
    del o, dm, df, dz
    torch.cuda.empty_cache()
    # C4: unmet + decision, 30-day backlog, now inside a European DST day
    n4 = 10_000_000
    m4 = synth.gen_specs(n4, seed=41, dense_minutes=0.5)
    z4 = synth.gen_zone_ids(n4, len(names), dst, seed=41)
    now = synth.ts(2026, 3, 29, 12)
    e4 = now - 30 * 86400 + synth.gen_times(n4, -86400, 86400, seed=41)
    rng = np.random.default_rng(41)
    pol4 = rng.integers(0, 3, n4).astype(np.uint8)
    act4 = rng.choice([0, 1, 2], size=n4, p=[.6, .3, .1]).astype(np.uint32)
    fl4 = (rng.random(n4) < 0.02).astype(np.uint8)
    dl4 = np.where(rng.random(n4) < 0.01, now - 100, -2**63).astype(np.int64)
    t = lambda a: torch.from_numpy(a).cuda()  # noqa: E731
    dm4, dz4, de4, dp, da, dfl, ddl = t(m4.view(np.int64)), t(z4), t(e4), t(pol4), t(act4), t(fl4), t(dl4)
    o4 = dict(last=torch.empty(n4, dtype=torch.int64, device="cuda"), next=torch.empty(n4, dtype=torch.int64, device="cuda"),
              missed=torch.empty(n4, dtype=torch.int32, device="cuda"), status=torch.empty(n4, dtype=torch.uint8, device="cuda"),
              decision=torch.empty(n4, dtype=torch.uint8, device="cuda"))
    # C4's specs are 5-field (C2's distribution): no seconds plane is passed (second 0, as ParseStandard yields)
    s = timed(torch, lambda: ctx.unmet_since_dev(dm4, de4, now, o4, 0, dp, da, dfl, ddl, dz4, stream=st, five_field=True), flush=flush)
    ctx.sync_check(stream=st)
    missed_total = int(o4["missed"].to(torch.int64).sum().item())
    ipc = NCU.get("unmet_warp_inst_per_cron")
    c4 = {"specs": n4, "specs_per_s": n4 / s, "ms": s * 1e3, "bytes_per_spec": 88,
          "roofline_frac": 88 * n4 / s / 1e9 / peak,
          "issue_frac": issue_frac(ipc, n4, s, clocks) if ipc else None, "warp_inst_per_cron": ipc,
          "issue_source": NCU.get("unmet_source"), "missed_runs_counted": missed_total,
          "equivalent_literal_next_calls_per_s": (missed_total + 2 * n4) / s,
          "note": "missed runs are counted analytically; not literal Next() calls"}
    dev_out = {k_: v.cpu().numpy() for k_, v in o4.items()}
    del dm4, dz4, de4, o4, dp, da, dfl, ddl
    torch.cuda.empty_cache()
    # C4 end to end: b200cron_unmet_since on pinned host buffers -- the call a batching reconciler makes
    # (cron_controller.go:184) -- and through a resident spec set
    P = api.PinnedArray
    hm, hz, he, hp, ha, hf, hd = (P((6, n4), np.uint64), P(n4, np.int32), P(n4, np.int64), P(n4, np.uint8),
                                  P(n4, np.uint32), P(n4, np.uint8), P(n4, np.int64))
    hm.array[:] = m4; hz.array[:] = z4; he.array[:] = e4; hp.array[:] = pol4; ha.array[:] = act4; hf.array[:] = fl4; hd.array[:] = dl4
    owners = dict(last=P(n4, np.int64), next=P(n4, np.int64), missed=P(n4, np.uint32), status=P(n4, np.uint8), decision=P(n4, np.uint8))
    ho = {k_: v.array for k_, v in owners.items()}
    call = lambda: ctx.unmet_since(hm.array, he.array, now, 0, hp.array, ha.array, hf.array, hd.array, hz.array, out=ho, five_field=True)  # noqa: E731
    call()
    best = 1e30
    for _ in range(3):
        t0 = time.perf_counter()
        call()
        best = min(best, time.perf_counter() - t0)
    same = all(np.array_equal(ho[k_], dev_out[k_].view(ho[k_].dtype)) for k_ in ho)
    c4["e2e"] = {"specs_per_s": n4 / best, "ms": best * 1e3, "h2d_bytes": 66 * n4, "d2h_bytes": 22 * n4,
                 "matches_device_path": bool(same),
                 "how": "b200cron_unmet_since on pinned host buffers (staged chunk copies + kernels + copies back), wall clock"}
    sset = ctx.specset(m4, z4, five_field=True)
    call = lambda: sset.unmet_since(he.array, now, 0, hp.array, ha.array, hf.array, hd.array, out=ho)  # noqa: E731
    call()
    best = 1e30
    for _ in range(3):
        t0 = time.perf_counter()
        call()
        best = min(best, time.perf_counter() - t0)
    same = all(np.array_equal(ho[k_], dev_out[k_].view(ho[k_].dtype)) for k_ in ho)
    sset.close()
    c4["e2e_resident"] = {"specs_per_s": n4 / best, "ms": best * 1e3, "h2d_bytes": 22 * n4, "d2h_bytes": 22 * n4,
                          "matches_device_path": bool(same),
                          "how": "b200cron_unmet_since_set: masks + zone ids resident in HBM, per-tick arrays pinned"}
    # CPU arm for C4: the oracle's literal chain on a bounded sample (every 2500th Cron), all host threads
    from oracle import pyoracle as po
    idx = np.arange(0, n4, 2500)
    ms = np.ascontiguousarray(m4[:, idx])
    t0 = time.perf_counter()
    w = po.batch_unmet(ms, e4[idx], now, 0, pol4[idx], act4[idx], fl4[idx], dl4[idx], z4[idx], names)
    dt = time.perf_counter() - t0
    ok = all(np.array_equal(w[k_], dev_out[k_][idx].view(w[k_].dtype)) for k_ in w)
    c4["cpu_baseline"] = {"specs_per_s": len(idx) / dt, "cores": po.cpu_count(), "kind": "port",
                          "sample": f"every 2500th Cron of the 10M ({len(idx)} Crons), literal Next() chain",
                          "sample_parity_vs_gpu": bool(ok)}
    out["C4_unmet_since"] = c4
    return out


def _c5_local(ctx, torch, st, rank, world, flush):
    """this rank's shard of C5; returns (n, eval seconds, history seconds, records, chunks, parity)"""
    from oracle import pyoracle as po
    N5 = 1_000_000_000
    lo, hi = N5 * rank // world, N5 * (rank + 1) // world
    n = hi - lo
    # ---- evaluation part: next fire of every spec of this rank's shard
    dm = torch.empty((6, n), dtype=torch.int64, device="cuda")
    df = torch.empty(n, dtype=torch.int64, device="cuda")
    ctx.synth_specs_dev(dm, n, 0xC0FFEE, lo, False, stream=st)
    ctx.synth_times_dev(df, n, T_LO, T_HI, 0xC0FFEE, lo, stream=st)
    o = torch.empty((n, 1), dtype=torch.int64, device="cuda")
    s_eval = timed(torch, lambda: ctx.next_n_dev(dm, df, o, 1, None, stream=st, five_field=True), reps=3, flush=flush)
    ctx.sync_check(stream=st)
    idx = np.arange(lo + (7 * rank) % 99_991, hi, 99_991, dtype=np.uint64)
    want = po.batch_next_n(synth.gen_specs(0, 0xC0FFEE, idx=idx), synth.gen_times(0, T_LO, T_HI, 0xC0FFEE, idx=idx), 1)
    got = o[torch.from_numpy((idx - np.uint64(lo)).astype(np.int64)).cuda()].cpu().numpy()
    parity = bool(np.array_equal(got, want))
    del dm, df, o
    torch.cuda.empty_cache()
    # ---- history part: one segment per Cron, U{0..16} records, keys = creation seconds within one day (ties);
    # as many segments at a time as fit beside their records (13 B per record + 24 B per segment of inputs)
    free = torch.cuda.mem_get_info()[0]
    per_seg = 24 + 8 * (8 + 4 + 1) + 16
    chunk = int(min(n, max(1_000_000, (free * 0.6) // per_seg)))
    s_hist, recs, chunks = 0.0, 0, 0
    base = 1_760_000_000
    c0 = lo
    while c0 < hi:
        cn = min(chunk, hi - c0)
        lens = torch.empty(cn, dtype=torch.int64, device="cuda")
        ctx.synth_times_dev(lens, cn, 0, 17, 0x515, c0, stream_id=301, stream=st)
        ctx.sync_check(stream=st)
        seg = torch.zeros(cn + 1, dtype=torch.int64, device="cuda")
        torch.cumsum(lens, 0, out=seg[1:])
        m = int(seg[-1].item())
        del lens
        key = torch.empty(max(m, 1), dtype=torch.int64, device="cuda")
        ctx.synth_times_dev(key, m, base - 86399, base + 1, 0x515 + c0 % 1_000_003, 0, stream_id=302, stream=st)
        lim = torch.full((cn,), 10, dtype=torch.int64, device="cuda")
        perm = torch.empty(max(m, 1), dtype=torch.int32, device="cuda")
        keep = torch.empty(max(m, 1), dtype=torch.uint8, device="cuda")
        s_hist += timed(torch, lambda: ctx.history_prune_dev(seg, key[:m], lim, perm[:m], keep[:m], stream=st), reps=2, flush=flush)
        ctx.sync_check(stream=st)
        recs += m
        chunks += 1
        sidx = torch.arange(0, cn, 9_973, device="cuda")[:200]  # sampled parity: every 9,973rd segment of the chunk
        a, b = seg[sidx].cpu().numpy(), seg[sidx + 1].cpu().numpy()
        for x, y in zip(a, b):
            wp, wk = po.history_prune(key[x:y].cpu().numpy(), 10)
            parity = (parity and np.array_equal(perm[x:y].cpu().numpy().view(np.uint32), wp)
                      and np.array_equal(keep[x:y].cpu().numpy(), wk))
        del seg, key, lim, perm, keep
        torch.cuda.empty_cache()
        c0 += cn
    return n, s_eval, s_hist, recs, chunks, parity


def c5_sharded(ctx, torch, dist, st, peak, rank, world, flush):
    """BASELINE config 5: 1e9 specs + finished-history sort/prune (historyLimit 10), sharded contiguously over the
    ranks with no collective on the data path.  Everything is generated on the device from the global index;
    parity is checked on a sample per rank against the oracle.  A rank that fails still joins the reductions."""
    from cron_operator_b200 import shard
    N5 = 1_000_000_000
    err = None
    try:
        n, s_eval, s_hist, recs, chunks, parity = _c5_local(ctx, torch, st, rank, world, flush)
    except Exception as e:
        err, n, s_eval, s_hist, recs, chunks, parity = repr(e), N5 // world, float("inf"), float("inf"), 0, 0, False
    s_eval, s_hist = shard.max_over_ranks([s_eval, s_hist], device="cuda")
    recs_all, par_all = recs, parity
    if world > 1:
        tt = torch.tensor([recs, int(parity)], dtype=torch.int64, device="cuda")
        dist.all_reduce(tt[:1], op=dist.ReduceOp.SUM)
        dist.all_reduce(tt[1:], op=dist.ReduceOp.MIN)
        recs_all, par_all = int(tt[0].item()), bool(tt[1].item())
    if err or s_eval == float("inf"):
        return {"error": err or "another rank failed"}
    return {"specs_total": N5, "specs_per_gpu": n, "evals_per_s": N5 / s_eval, "eval_ms": s_eval * 1e3,
            "eval_bytes_per_spec": 56, "eval_roofline_frac": 56 * n / s_eval / 1e9 / peak,  # no seconds plane: 40 + 8 + 8
            "history_records": recs_all, "records_per_s": recs_all / s_hist, "history_ms": s_hist * 1e3,
            "history_roofline_frac": (13 * recs + 16 * n) / s_hist / 1e9 / peak, "history_chunks_per_gpu": chunks,
            "total_ms": (s_eval + s_hist) * 1e3, "sampled_parity": par_all,
            "parity_sample": "every 99,991st spec and up to 200 of every 9,973rd segment per chunk, per rank, vs the oracle",
            "timing": "device time per rank (CUDA events, L2 flushed before each launch), max over ranks"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-extra", action="store_true", help="skip the C3/C4/C5 measurements")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--no-overlap", action="store_true",
                    help="do not let back-to-back launches overlap (b200cron_ctx_set_overlap off)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from cron_operator_b200 import api, shard

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = api.Context(devices=[local])
    # a dedicated (non-default) torch stream: the ABI reads stream handle 0 as "the context's own stream", and the
    # CUDA events below must sit on the stream the kernels are launched on
    tstream = torch.cuda.Stream()
    torch.cuda.set_stream(tstream)
    st = tstream.cuda_stream
    assert st != 0
    peak, peak_src = peaks()

    # This rank's contiguous shard of the synthetic spec list: N_BATCH distinct C2 batches of 1M specs each, generated
    # on the device by global index (bit-identical to synth.py).  Launches rotate through the batches, so a batch's
    # 64 MB were last touched N_BATCH-1 launches (448 MB of other traffic) ago and are gone from the 126 MB L2.
    base = rank * N_BATCH * N_SPECS
    d_masks, d_from, d_out, planes, planes5 = [], [], [], [], []
    for b in range(N_BATCH):
        dm = torch.empty((6, N_SPECS), dtype=torch.int64, device="cuda")
        df = torch.empty(N_SPECS, dtype=torch.int64, device="cuda")
        ctx.synth_specs_dev(dm, N_SPECS, 0xC0FFEE, base + b * N_SPECS, False, stream=st)
        ctx.synth_times_dev(df, N_SPECS, T_LO, T_HI, 0xC0FFEE, base + b * N_SPECS, stream=st)
        d_masks.append(dm)
        d_from.append(df)
        d_out.append(torch.empty((N_SPECS, 1), dtype=torch.int64, device="cuda"))
        planes.append(api._planes(dm)[0])                    # all six planes: the 64 B/eval of SURVEY.md 8d really move
        planes5.append(api._planes(dm, five_field=True)[0])  # the same specs as a 5-field caller passes them (no seconds plane)
    ctx.sync_check(stream=st)
    masks = synth.gen_specs(N_SPECS, start=base)
    frm = synth.gen_times(N_SPECS, T_LO, T_HI, start=base)
    assert np.array_equal(d_masks[0].cpu().numpy().view(np.uint64), masks)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    L = api.lib()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def launch(b, pl=None):  # the C-ABI call, with the argument marshalling hoisted out of the timed loop
        rc = L.b200cron_next_n_dev(ctx._h, 0, st, (pl or planes)[b], None, d_from[b].data_ptr(), N_SPECS, 1, d_out[b].data_ptr())
        if rc != 0:
            ctx._check(rc)

    def step(i):
        for j in range(STEP_BATCHES):
            launch((i * STEP_BATCHES + j) % N_BATCH)

    # Launches are independent batches with their own input and output buffers, so a launch may start on SMs the
    # previous one has left idle (programmatic dependent launch, see b200cron_ctx_set_overlap in the header).
    ctx.set_overlap(not args.no_overlap)
    for i in range(args.warmup):
        step(i)
    ctx.sync_check(stream=st)

    # ---- kernel-resident timing: K steps between one pair of CUDA events on the launch stream ----
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = ctx.launch_count
    barrier()
    e0.record()
    for i in range(args.steps):
        step(i)
    e1.record()
    barrier()
    launches = ctx.launch_count - launches0
    ctx.sync_check(stream=st)
    dev_s = e0.elapsed_time(e1) * 1e-3
    # for the record: the same launches with every launch waiting for the previous one to finish (overlap off)
    serial_us = None
    if not args.no_overlap:
        ctx.set_overlap(False)
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record()
        for i in range(min(args.steps, 5)):
            step(i)
        s1.record()
        ctx.sync_check(stream=st)
        serial_us = s0.elapsed_time(s1) / (min(args.steps, 5) * STEP_BATCHES) * 1e3
        ctx.set_overlap(True)
    # the same launches as a 5-field caller makes them: no seconds plane (second 0 is all ParseStandard can produce), so
    # 56 B/eval move and the kernel's 5-field instantiation runs
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for j in range(N_BATCH):
        launch(j, planes5)
    f0.record()
    for j in range(5 * STEP_BATCHES):
        launch(j % N_BATCH, planes5)
    f1.record()
    ctx.sync_check(stream=st)
    five_us = f0.elapsed_time(f1) / (5 * STEP_BATCHES) * 1e3
    # cross-check: the same kernel timed one launch at a time with the L2 flushed before each (includes launch latency)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(30)]
    for i, (a, b) in enumerate(ev):
        flush.zero_()
        a.record()
        launch(i % N_BATCH)
        b.record()
    torch.cuda.synchronize()
    kernel_us = [a.elapsed_time(b) * 1e3 for a, b in ev]
    launch(0)
    ctx.sync_check(stream=st)
    d_out0 = d_out[0].cpu().numpy()

    # ---- end to end through the host-pointer C ABI (pinned host buffers) ----
    h_masks, h_from, h_out = api.PinnedArray((6, N_SPECS), np.uint64), api.PinnedArray(N_SPECS, np.int64), api.PinnedArray((N_SPECS, 1), np.int64)
    h_masks.array[:] = masks
    h_from.array[:] = frm
    e2e_calls = max(50, args.steps)
    # 5-field schedules (the reference's only kind) always fire at second 0: the seconds plane is not sent (NULL)
    for _ in range(5):
        ctx.next_n(h_masks.array, h_from.array, 1, out=h_out.array, five_field=True)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_calls):
        ctx.next_n(h_masks.array, h_from.array, 1, out=h_out.array, five_field=True)
    e2e_s = time.perf_counter() - t0
    barrier()
    assert np.array_equal(h_out.array, d_out0), "host-pointer and device-pointer paths disagree"
    # the same through a resident spec set: the immutable masks stay in HBM, a call moves 8 B in and 8 B out per spec
    sset = ctx.specset(masks, None, five_field=True)
    h_out.array[:] = 0
    for _ in range(5):
        sset.next_n(h_from.array, 1, out=h_out.array)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_calls):
        sset.next_n(h_from.array, 1, out=h_out.array)
    e2e_set_s = time.perf_counter() - t0
    barrier()
    assert np.array_equal(h_out.array, d_out0), "spec-set and device-pointer paths disagree"
    sset.close()
    clocks = sampler.stop() if rank == 0 else None

    dev_s, e2e_s, e2e_set_s = shard.max_over_ranks([dev_s, e2e_s, e2e_set_s], device="cuda")  # identity at N=1

    line = None
    if rank == 0:
        n_launch = args.steps * STEP_BATCHES
        value = world * N_SPECS * n_launch / dev_s
        avg_launch_s = dev_s / n_launch
        achieved = BYTES_PER_EVAL * N_SPECS / avg_launch_s / 1e9
        ipe = NCU.get("next1_warp_inst_per_eval")
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_s / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "specs_per_launch_per_gpu": N_SPECS, "launches_per_step": STEP_BATCHES,
                       "specs_per_step_per_gpu": N_SPECS * STEP_BATCHES, "k": 1, "zone": "UTC",
                       "l2": "inputs larger than L2: launches rotate through 8 distinct 1M-spec batches (512 MB working "
                             "set per GPU), so no launch finds its batch in the 126 MB L2",
                       "timing": "one CUDA-event pair on the launch stream around the K steps, max over ranks",
                       "overlap": ("off" if args.no_overlap else "on: the batches are independent, so the library "
                                   "launches them with programmatic stream serialization -- a launch's thread blocks "
                                   "fill SMs while the previous launch drains; every kernel executes "
                                   "griddepcontrol.wait before its first store, so completion stays in stream order"),
                       "parallelism": f"specs sharded contiguously over {world} GPU(s), no collective"},
            "e2e": {"value": world * N_SPECS * e2e_calls / e2e_s, "unit": UNIT, "h2d_bytes_per_step": 48 * N_SPECS,
                    "d2h_bytes_per_step": 8 * N_SPECS, "ms_per_call": e2e_s / e2e_calls * 1e3, "calls": e2e_calls,
                    "how": "b200cron_next_n on pinned host buffers, 1M specs per call, wall clock per call (it returns "
                           "after the results are in host memory): the kernel reads the 5 mask planes + from and writes "
                           "the results in place over PCIe (seconds plane NULL = second 0, as ParseStandard yields)",
                    "resident": {"value": world * N_SPECS * e2e_calls / e2e_set_s, "unit": UNIT,
                                 "h2d_bytes_per_step": 8 * N_SPECS, "d2h_bytes_per_step": 8 * N_SPECS,
                                 "ms_per_call": e2e_set_s / e2e_calls * 1e3,
                                 "how": "b200cron_next_n_set: the spec set (immutable after ParseStandard) is resident "
                                        "in HBM; per call only `from` goes in and `out` comes back (pinned, in place)"}},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": NCU.get("next1_dram_bytes_per_launch"), "traffic_source": NCU.get("next1_source"),
                         "peak_source": peak_src, "kernel": NCU.get("next1_kernel", "next1_ring_kernel<true>"),
                         "algorithmic_bytes_per_launch": BYTES_PER_EVAL * N_SPECS,
                         "us_per_launch": avg_launch_s * 1e6, "us_per_launch_overlap_off": serial_us,
                         "frac_overlap_off": (BYTES_PER_EVAL * N_SPECS / (serial_us * 1e-6) / 1e9 / peak) if serial_us else None,
                         "issue_frac": issue_frac(ipe, N_SPECS, avg_launch_s, clocks) if ipe else None,
                         "warp_inst_per_eval": ipe,
                         "five_field_launch": {"us_per_launch": five_us, "evals_per_s_per_gpu": N_SPECS / (five_us * 1e-6),
                                               "bytes_per_eval": 56, "frac": 56 * N_SPECS / (five_us * 1e-6) / 1e9 / peak,
                                               "note": "same specs without the seconds plane (NULL = second 0), 5-field "
                                                       "instantiation; this GPU only, not max over ranks"},
                         "flushed_single_launch_us_min_med_max": [min(kernel_us), float(np.median(kernel_us)), max(kernel_us)],
                         "note": "cross-check: one launch at a time after a 256 MiB L2 flush, includes launch latency"},
            "clocks": clocks,
        }
        if not args.no_cpu:
            _, ms1, _ = cpu_reference(2, 1)                      # calibrate: ~12 s of wall time on all host threads
            reps = int(min(400, max(20, 12_000 / max(ms1, 1e-3))))
            v, ms, threads = cpu_reference(reps, 1)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": f"full C2 batch ({N_SPECS} specs) x {reps} reps, all host threads; C "
                                              "restatement of robfig/cron v3.0.1 + Go time (no Go toolchain here)"}
    if not args.no_extra:
        extra = {}
        if world > 1:
            dist.barrier()
        try:
            extra["C5_sharded"] = c5_sharded(ctx, torch, dist, st, peak, rank, world, flush)
        except Exception as e:  # the headline must survive a failure of the side measurements
            extra["C5_sharded"] = {"error": repr(e)}
        if rank == 0 and world == 1:
            try:
                extra.update(c3_c4(ctx, torch, st, peak, flush, clocks))
            except Exception as e:
                extra["C3_C4"] = {"error": repr(e)}
        if rank == 0:
            line["extra"] = extra
    if rank == 0:
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
