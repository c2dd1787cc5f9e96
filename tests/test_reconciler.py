"""Host logic of the batching reconciler (cron_operator_b200/reconciler.py): one tick over
a mixed bag of Crons must report, per Cron, exactly what the reference's Reconcile would do
(internal/controller/cron_controller.go:163-238, :307-346).  On CPU the engine is a checker built on the oracle; the
GPU test runs the same tick through the CUDA library and expects the same results."""
import numpy as np
import pytest

from cron_operator_b200 import api, synth
from cron_operator_b200.reconciler import BatchReconciler, CronView
from oracle import pyoracle as po

ts = synth.ts
NOW = ts(2026, 1, 1, 12)


class OracleEngine:
    """the two batched calls of api.Context, answered by the CPU oracle (test infrastructure)"""

    def __init__(self):
        from tests import shim
        self.names = shim.zone_names()

    def UnmetSince(self, specs, earliest, now, policy=None, active=None, flags=None, deadline=None, now_nsec=0):
        zid = np.array([self.names.index(s.Location) if s.Location else 0 for s in specs], dtype=np.int32)
        pol = None if policy is None else np.asarray(policy, dtype=np.uint8)
        return po.batch_unmet(api.pack_specs(specs), np.asarray(earliest, dtype=np.int64), now, now_nsec, pol, active,
                              flags, deadline, zid, self.names)

    def history_prune(self, seg_off, key, limit=None):
        return po.batch_history(np.asarray(seg_off, dtype=np.uint64), np.asarray(key, dtype=np.int64), limit)


def _crons():
    return [
        CronView("a", "*/1 * * * *", NOW - 300),                                                    # K3: 5 missed
        CronView("b", "*/1 * * * *", NOW - 300, "Forbid", active=2),
        CronView("c", "*/1 * * * *", NOW - 300, "Replace", active=1),
        CronView("d", "*/1 * * * *", NOW - 300, suspend=True),
        CronView("e", "not a schedule", NOW - 300),                                                 # K1
        CronView("f", "0 0 30 2 *", NOW - 300),                                                     # K2
        CronView("g", "@every 90s", NOW - 200),
        CronView("h", "*/1 * * * *", NOW - 300, deadline=NOW - 1),
        CronView("i", "*/5 * * * *", NOW - 86400, last_schedule_time=NOW),
        CronView("j", "TZ=Asia/Tokyo 0 9 * * *", NOW - 2 * 86400),
        CronView("k", "*/5 * * * *", NOW - 60, history_limit=3, finished_created_at=[5, 3, 9, 3, 7]),
        CronView("l", "not a schedule", NOW - 300, suspend=True),
        CronView("m", "* * * * *", NOW - 7200),
        CronView("n", "*/5 * * * *", NOW - 60, deletion_timestamp=NOW - 5, history_limit=0, finished_created_at=[2, 1]),
        CronView("o", "*/5 * * * *", NOW + 30, finished_created_at=[4, 4, 1]),                      # created "after now"
        CronView("p", "@every 1h", NOW - 100, "Forbid", active=1),
    ]


def _check(res):
    r = {x.cron: x for x in res}
    assert (r["a"].decision, r["a"].missed, r["a"].requeue_after) == (api.DEC_CREATE, 5, 60)
    assert r["a"].create == f"a-{NOW + 60}" and r["a"].set_last_schedule_time == NOW and not r["a"].delete_active
    assert (r["b"].decision, r["b"].requeue_after, r["b"].create, r["b"].set_last_schedule_time) == (api.DEC_SKIP_FORBID, 60, None, None)
    assert r["c"].decision == api.DEC_REPLACE_CREATE and r["c"].delete_active and r["c"].create == f"c-{NOW + 60}"
    assert (r["d"].decision, r["d"].requeue_after, r["d"].create, r["d"].error) == (api.DEC_GATE_SUSPENDED, None, None, None)
    assert r["e"].error.startswith('unparsable cron "not a schedule"') and r["e"].requeue_after is None and r["e"].create is None
    assert r["f"].error == 'unschedulable cron "0 0 30 2 *"' and r["f"].requeue_after is None
    assert (r["g"].decision, r["g"].missed, r["g"].requeue_after, r["g"].create) == (api.DEC_CREATE, 2, 90, f"g-{NOW + 90}")
    assert r["h"].decision == api.DEC_GATE_DEADLINE and r["h"].deadline_event and r["h"].requeue_after is None
    assert (r["i"].decision, r["i"].missed, r["i"].requeue_after, r["i"].create) == (api.DEC_REQUEUE_ONLY, 0, 300, None)
    assert (r["j"].decision, r["j"].missed, r["j"].requeue_after) == (api.DEC_CREATE, 2, 43200)
    assert r["j"].create == f"j-{ts(2026, 1, 2)}"
    assert (r["k"].history_delete, r["k"].history_keep) == ([1, 3], [0, 4, 2])
    assert (r["l"].decision, r["l"].error) == (api.DEC_GATE_SUSPENDED, None)
    assert r["m"].missed == 120 and r["m"].too_many_missed and r["m"].decision == api.DEC_CREATE
    assert r["n"].decision == api.DEC_GATE_DELETED and r["n"].history_delete == [1, 0] and r["n"].history_keep == []
    assert (r["o"].decision, r["o"].requeue_after, r["o"].history_keep) == (api.DEC_REQUEUE_ONLY, 300, [2, 0, 1])
    assert (r["p"].decision, r["p"].requeue_after, r["p"].missed) == (api.DEC_REQUEUE_ONLY, 3600, 0)
    assert [x.cron for x in res] == [c.name for c in _crons()]


def test_tick_reports_what_reconcile_would_do():
    rec = BatchReconciler(OracleEngine())
    _check(rec.reconcile_tick(_crons(), NOW))
    # a sub-second "now" is after a deadline of the same second (now.After(deadline), :176)
    c = [CronView("x", "*/1 * * * *", NOW - 300, deadline=NOW), CronView("y", "@every 30s", NOW - 300, deadline=NOW)]
    assert [r.decision for r in rec.reconcile_tick(c, NOW, 0)] == [api.DEC_CREATE, api.DEC_CREATE]
    assert [r.decision for r in rec.reconcile_tick(c, NOW, 1)] == [api.DEC_GATE_DEADLINE, api.DEC_GATE_DEADLINE]
    assert rec.reconcile_tick([], NOW) == []


def test_tick_is_one_batched_call_per_kind():
    class Counting(OracleEngine):
        calls = {"unmet": 0, "history": 0}

        def UnmetSince(self, *a, **k):
            self.calls["unmet"] += 1
            return super().UnmetSince(*a, **k)

        def history_prune(self, *a, **k):
            self.calls["history"] += 1
            return super().history_prune(*a, **k)

    eng = Counting()
    BatchReconciler(eng).reconcile_tick(_crons(), NOW)
    assert eng.calls == {"unmet": 1, "history": 1}


@pytest.mark.gpu
def test_tick_through_the_cuda_library():
    ctx = api.Context()
    try:
        _check(BatchReconciler(ctx).reconcile_tick(_crons(), NOW))
    finally:
        ctx.close()
