#!/usr/bin/env python
"""bench.py -- schedule evaluations/sec of the batched cron-schedule path on B200 (BASELINE.json metric).

A "step" is one pass of the hot path over one batch of synthetic specs.  Workload at every N: BASELINE config C2
(1M random 5-field UTC cron specs, next single fire time) per GPU (weak scaling: each rank owns its own contiguous
1M-spec shard of one big synthetic spec list; no collective on the data path).

  value     evals/s with inputs resident in HBM: K back-to-back steps between one CUDA-event pair on the launch
            stream; steps rotate through 8 distinct 1M-spec batches (512 MB > L2), max over ranks
  e2e       the same metric through the reference-facing C ABI with pinned HOST buffers (H2D + kernel + D2H inside)
  roofline  algorithmic bytes (64 B/eval) / kernel time vs MEASURED_PEAKS.json hbm_gbs
  cpu_baseline  the C restatement of the reference's Go path (oracle/) on this box's host cores (rank 0, N=1)

--impl reference times that CPU path alone (the reference is Go; no Go toolchain exists on this image, so the
reference arm is the oracle port, kind "port").
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
BYTES_PER_EVAL = 64          # 48 masks + 8 from + 8 out
T_LO, T_HI = synth.ts(2024), synth.ts(2034)
NCU_TRAFFIC_BYTES = 56_009_216 + 386_560  # one launch of next1_kernel, profiles/r1h_ncu_next1_summary.md
METRIC = "schedule_evals_per_sec"
UNIT = "evals/s"
WORKLOAD = "C2: 1M random 5-field UTC cron specs, next single fire time (per GPU)"


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.rows, self.p = gpu_index, [], None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "200", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
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
        "config": {"workload": WORKLOAD, "note": "CPU path; every step is the full 1M-spec C2 batch on host cores"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"full C2 batch ({N_SPECS} specs) x {args.steps} steps; C restatement of "
                                   "robfig/cron v3.0.1 + Go time (no Go toolchain on this image)"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def extra_kernels(ctx, torch, st, peak):
    """kernel-only numbers for the other BASELINE configs (bounded sizes), each with its roofline fraction"""
    out = {}
    names = ctx.zone_names
    dst = np.array([z for z in range(1, len(names)) if (ctx.zone_periods(z)[0] > synth.ts(2030)).any()])

    def timed(fn, reps=5):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e30
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        for _ in range(reps):
            flush.zero_()
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return best * 1e-3

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
    s = timed(lambda: ctx.next_n_dev(dm, df, o, k, dz, stream=st), reps=3)
    out["C3_nextN_k16_zoned_100M"] = {"specs": n, "evals_per_s": n * k / s, "specs_per_s": n / s, "ms": s * 1e3,
                                      "bytes_per_spec": 188, "roofline_frac": 188 * n / s / 1e9 / peak,
                                      "data": "generated on device (b200cron_synth_*_dev), seed 3"}
    del o
    # C5 evaluation part on one GPU: 1B UTC specs, next fire (64 GB resident) when the device has the room
    try:
        if torch.cuda.mem_get_info()[0] > 90 * (1 << 30):
            del dm, df, dz
            n5 = 1_000_000_000
            dm5 = torch.empty((6, n5), dtype=torch.int64, device="cuda")
            df5 = torch.empty(n5, dtype=torch.int64, device="cuda")
            ctx.synth_specs_dev(dm5, n5, 0xC0FFEE, 0, False, stream=st)
            ctx.synth_times_dev(df5, n5, T_LO, T_HI, 0xC0FFEE, 0, stream=st)
            o5 = torch.empty((n5, 1), dtype=torch.int64, device="cuda")
            s = timed(lambda: ctx.next_n_dev(dm5, df5, o5, 1, None, stream=st), reps=3)
            out["C5_eval_part_1B_specs"] = {"specs": n5, "evals_per_s": n5 / s, "ms": s * 1e3,
                                            "roofline_frac": 64 * n5 / s / 1e9 / peak}
            del dm5, df5, o5
    except Exception as e:
        out["C5_eval_part_1B_specs"] = {"error": repr(e)}
    torch.cuda.empty_cache()
    # C4 shape: unmet + decision, 30-day backlog, now inside a European DST day
    n4 = 10_000_000
    m4 = synth.gen_specs(n4, seed=41, dense_minutes=0.5)
    z4 = synth.gen_zone_ids(n4, len(names), dst, seed=41)
    now = synth.ts(2026, 3, 29, 12)
    e4 = now - 30 * 86400 + synth.gen_times(n4, -86400, 86400, seed=41)
    rng = np.random.default_rng(41)
    t = lambda a: torch.from_numpy(a).cuda()  # noqa: E731
    dm4, dz4, de4 = t(m4.view(np.int64)), t(z4), t(e4)
    dp, da = t(rng.integers(0, 3, n4).astype(np.uint8)), t(rng.choice([0, 1, 2], size=n4, p=[.6, .3, .1]).astype(np.uint32))
    dfl, ddl = t((rng.random(n4) < 0.02).astype(np.uint8)), t(np.where(rng.random(n4) < 0.01, now - 100, -2**63).astype(np.int64))
    o4 = dict(last=torch.empty(n4, dtype=torch.int64, device="cuda"), next=torch.empty(n4, dtype=torch.int64, device="cuda"),
              missed=torch.empty(n4, dtype=torch.int32, device="cuda"), status=torch.empty(n4, dtype=torch.uint8, device="cuda"),
              decision=torch.empty(n4, dtype=torch.uint8, device="cuda"))
    s = timed(lambda: ctx.unmet_since_dev(dm4, de4, now, o4, 0, dp, da, dfl, ddl, dz4, stream=st))
    missed_total = int(o4["missed"].to(torch.int64).sum().item())
    out["C4_unmet_since"] = {"specs": n4, "specs_per_s": n4 / s, "ms": s * 1e3, "bytes_per_spec": 96,
                             "roofline_frac": 96 * n4 / s / 1e9 / peak, "missed_runs_counted": missed_total,
                             "equivalent_literal_next_calls_per_s": (missed_total + 2 * n4) / s,
                             "note": "missed runs are counted analytically; not literal Next() calls"}
    del dm4, dz4, de4, o4
    # C5 shape: history sort/prune, U{0..16} records per Cron, historyLimit 10 (8M segments)
    ns = 8_000_000
    lens = rng.integers(0, 17, ns)
    seg = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    m = int(seg[-1])
    key = (1_760_000_000 - rng.integers(0, 86400, m)).astype(np.int64)
    ds, dk, dl = t(seg.view(np.int64)), t(key), t(np.full(ns, 10, np.int64))
    perm, keep = torch.empty(m, dtype=torch.int32, device="cuda"), torch.empty(m, dtype=torch.uint8, device="cuda")
    s = timed(lambda: ctx.history_prune_dev(ds, dk, dl, perm, keep, stream=st))
    out["C5_history_prune"] = {"segments": ns, "records": m, "records_per_s": m / s, "ms": s * 1e3,
                               "roofline_frac": (13 * m + 16 * ns) / s / 1e9 / peak}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-extra", action="store_true", help="skip the C3/C4/C5 kernel-only measurements")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--no-overlap", action="store_true",
                    help="do not let back-to-back steps overlap (b200cron_ctx_set_overlap off)")
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

    # This rank's contiguous shard of the synthetic spec list: N_BATCH distinct C2 batches of 1M specs each.  A step is
    # one pass over ONE batch; steps rotate through the batches, so a batch's 64 MB were last touched N_BATCH-1 steps
    # (448 MB of other traffic) ago and are long gone from the 126 MB L2 when it comes round again.
    N_BATCH = 8
    base = rank * N_BATCH * N_SPECS
    d_masks, d_from, d_out, planes = [], [], [], []
    for b in range(N_BATCH):
        m = synth.gen_specs(N_SPECS, start=base + b * N_SPECS)
        f = synth.gen_times(N_SPECS, T_LO, T_HI, start=base + b * N_SPECS)
        if b == 0:
            masks, frm = m, f
        d_masks.append(torch.from_numpy(m.view(np.int64)).cuda())
        d_from.append(torch.from_numpy(f).cuda())
        d_out.append(torch.empty((N_SPECS, 1), dtype=torch.int64, device="cuda"))
        planes.append(api._planes(d_masks[b])[0])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2, for the cross-check below
    L = api.lib()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(b):  # the C-ABI call, with the argument marshalling hoisted out of the timed loop
        rc = L.b200cron_next_n_dev(ctx._h, 0, st, planes[b], None, d_from[b].data_ptr(), N_SPECS, 1, d_out[b].data_ptr())
        if rc != 0:
            ctx._check(rc)

    # Steps are independent batches with their own input and output buffers, so a step may start on SMs the previous
    # step's last wave has left idle (programmatic dependent launch, see b200cron_ctx_set_overlap in the header).
    ctx.set_overlap(not args.no_overlap)
    for i in range(max(args.warmup, N_BATCH)):
        step(i % N_BATCH)
    ctx.sync_check(stream=st)

    # ---- kernel-resident timing: K back-to-back steps between one pair of CUDA events on the launch stream ----
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = ctx.launch_count
    barrier()
    e0.record()
    for i in range(args.steps):
        step(i % N_BATCH)
    e1.record()
    barrier()
    launches = ctx.launch_count - launches0
    ctx.sync_check(stream=st)
    dev_s = e0.elapsed_time(e1) * 1e-3
    # for the record: the same K steps with every launch waiting for the previous one to finish (overlap off)
    serial_ms = None
    if not args.no_overlap:
        ctx.set_overlap(False)
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record()
        for i in range(args.steps):
            step(i % N_BATCH)
        s1.record()
        ctx.sync_check(stream=st)
        serial_ms = s0.elapsed_time(s1) / args.steps
        ctx.set_overlap(True)
    # cross-check: the same kernel timed one launch at a time with the L2 flushed before each (includes launch latency)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(min(args.steps, 30))]
    for i, (a, b) in enumerate(ev):
        flush.zero_()
        a.record()
        step(i % N_BATCH)
        b.record()
    torch.cuda.synchronize()
    kernel_ms = [a.elapsed_time(b) for a, b in ev]
    d_out0 = d_out[0]

    # ---- end to end through the host-pointer C ABI (pinned host buffers) ----
    h_masks, h_from, h_out = api.PinnedArray((6, N_SPECS), np.uint64), api.PinnedArray(N_SPECS, np.int64), api.PinnedArray((N_SPECS, 1), np.int64)
    h_masks.array[:] = masks
    h_from.array[:] = frm
    # 5-field schedules (the reference's only kind) always fire at second 0: the seconds plane is not sent (NULL)
    for _ in range(args.warmup):
        ctx.next_n(h_masks.array, h_from.array, 1, out=h_out.array, five_field=True)
    e2e_steps = args.steps
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        ctx.next_n(h_masks.array, h_from.array, 1, out=h_out.array, five_field=True)
    e2e_s = time.perf_counter() - t0
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    step(0)
    ctx.sync_check(stream=st)
    assert np.array_equal(h_out.array, d_out0.cpu().numpy()), "host-pointer and device-pointer paths disagree"

    dev_s, e2e_s = shard.max_over_ranks([dev_s, e2e_s], device="cuda")  # max over ranks (identity at N=1)

    if rank == 0:
        value = world * N_SPECS * args.steps / dev_s
        e2e = world * N_SPECS * e2e_steps / e2e_s
        avg_kernel_s = dev_s / args.steps
        achieved = BYTES_PER_EVAL * N_SPECS / avg_kernel_s / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": avg_kernel_s * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "specs_per_step_per_gpu": N_SPECS, "k": 1, "zone": "UTC",
                       "l2": "inputs larger than L2: steps rotate through 8 distinct 1M-spec batches (512 MB working "
                             "set per GPU), so no step finds its batch in the 126 MB L2",
                       "timing": "one CUDA-event pair on the launch stream around the K back-to-back steps, max over ranks",
                       "overlap": ("off" if args.no_overlap else "on: the steps are independent batches, so the library "
                                   "launches them with programmatic stream serialization -- a step's thread blocks "
                                   "fill SMs while the previous step's last wave drains; ms_per_step = region / K"),
                       "parallelism": f"specs sharded contiguously over {world} GPU(s), no collective"},
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": 48 * N_SPECS, "d2h_bytes_per_step": 8 * N_SPECS,
                    "ms_per_step": e2e_s / e2e_steps * 1e3,
                    "how": "b200cron_next_n on pinned host buffers, wall clock per call (it returns after the results "
                           "are in host memory): the kernel reads the 5 mask planes + from and writes the results in "
                           "place over PCIe (seconds plane NULL = second 0, as ParseStandard always yields)"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": NCU_TRAFFIC_BYTES, "traffic_source": "profiles/r1h_ncu_next1_summary.md "
                         "(dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full; the 8 MB output "
                         "is still largely in L2 when the kernel ends)",
                         "peak_source": peak_src, "kernel": "next1_kernel",
                         "algorithmic_bytes_per_launch": BYTES_PER_EVAL * N_SPECS,
                         "ms_per_step_overlap_off": serial_ms,
                         "flushed_single_launch_ms_min_med_max": [min(kernel_ms), float(np.median(kernel_ms)), max(kernel_ms)],
                         "note": "cross-check: one launch at a time after a 256 MiB L2 flush, includes launch latency"},
            "clocks": clocks,
        }
        if world == 1 and not args.no_cpu:
            _, ms1, _ = cpu_reference(2, 1)                      # calibrate: ~12 s of wall time on all host threads
            reps = int(min(400, max(20, 12_000 / max(ms1, 1e-3))))
            v, ms, threads = cpu_reference(reps, 1)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": f"full C2 batch ({N_SPECS} specs) x {reps} reps, all host threads; C "
                                              "restatement of robfig/cron v3.0.1 + Go time (no Go toolchain here)"}
        if world == 1 and not args.no_extra:
            try:
                line["extra"] = extra_kernels(ctx, torch, st, peak)
            except Exception as e:  # the headline must survive a failure of the side measurements
                line["extra"] = {"error": repr(e)}
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
