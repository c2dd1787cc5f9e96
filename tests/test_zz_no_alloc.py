"""The one entry point that never allocates: b200cron_unmet_list_offsets_static.

It exists for hosts that cannot let this library call cudaMalloc while they are using it (the GPU's memory already
belongs to the host's own pool allocator, or the loop cannot take the device-wide synchronisation an allocation
causes).  The test drives it through the C ABI and holds it to the allocating entry point's output, bit for bit (the
tolerance every integer test in this suite uses), on two calls in a row that share the library's fixed scratch array.
Collected last on purpose: if it fails under `-x`, every other test has already run.
"""
import numpy as np
import pytest

from cron_operator_b200 import api

pytestmark = pytest.mark.gpu


def _inputs(torch, n, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    missed = torch.randint(0, 300, (n,), generator=g, device="cuda", dtype=torch.int32)
    decision = torch.randint(1, 9, (n,), generator=g, device="cuda", dtype=torch.int32).to(torch.uint8)
    return missed, decision


def test_static_scratch_offsets_match_allocating_entry_point_twice_in_a_row():
    import torch
    ctx = api.Context(devices=[0])
    try:
        stream = torch.cuda.Stream()
        with torch.cuda.stream(stream):
            st = stream.cuda_stream
            # two calls back to back; the second is smaller, so it runs over what the first left in the array
            cases = [(1_500_003, 11, 100), (70_001, 12, 7)]
            inputs = [(_inputs(torch, n, seed), n, cap) for n, seed, cap in cases]
            want = []
            for (missed, decision), n, cap in inputs:
                off = torch.empty(n + 1, dtype=torch.int64, device="cuda")
                ctx.unmet_list_offsets_dev(missed, decision, off, cap, stream=st)   # existing, allocating entry point
                ctx.sync_check(stream=st)
                want.append(off.cpu().numpy())
            outs = [torch.empty(n + 1, dtype=torch.int64, device="cuda") for _, n, _ in inputs]
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
            for ((missed, decision), n, cap), off in zip(inputs, outs):
                ctx.unmet_list_offsets_static(missed, decision, off, cap, stream=st)  # new: allocates nothing
            ctx.sync_check(stream=st)
            assert torch.cuda.mem_get_info()[0] >= free_before, "device memory was allocated during the calls"
            got = [o.cpu().numpy() for o in outs]
            for g, w, ((missed, decision), n, cap) in zip(got, want, inputs):
                assert np.array_equal(g, w), (n, cap)
                # and against the definition itself: exclusive prefix sum of min(missed, cap) over listed decisions
                cnt = np.where(decision.cpu().numpy() >= api.DEC_REQUEUE_ONLY, np.minimum(missed.cpu().numpy(), cap), 0)
                assert np.array_equal(g, np.concatenate([[0], np.cumsum(cnt.astype(np.int64))])), (n, cap)
                assert g[0] == 0 and np.all(np.diff(g) >= 0) and np.all(np.diff(g) <= cap)
            # n == 0 and the size limit
            off1 = torch.full((1,), -1, dtype=torch.int64, device="cuda")
            ctx.unmet_list_offsets_static(inputs[0][0][0], inputs[0][0][1], off1, 5, stream=st, n=0)
            ctx.sync_check(stream=st)
            assert int(off1.item()) == 0
            with pytest.raises(api.B200CronError) as e:
                ctx.unmet_list_offsets_static(inputs[0][0][0], inputs[0][0][1], outs[0], 5, stream=st, n=16_777_217)
            assert e.value.code == api.EINVAL and "16777216" in str(e.value)
    finally:
        ctx.close()
