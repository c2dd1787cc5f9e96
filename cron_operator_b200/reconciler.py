"""A batching reconciler over the schedule path.

The reference reconciles one Cron per call (internal/controller/cron_controller.go:90-238): parse the schedule, walk
sched.Next() from the last schedule time to now, decide, order and prune the history.  Every step of that which is
arithmetic is the same for all Crons, so a tick here takes the Crons that are due (what the controller-runtime work
queue would hand to up to MaxConcurrentReconciles goroutines, cmd/operator/start.go:174-176), makes ONE UnmetSince
call and ONE history_prune call for all of them, and fans the answers back as per-Cron results that say exactly what
Reconcile would do next: the ctrl.Result, the workloads to delete, the workload to create, the status fields to set.
The Kubernetes calls themselves (Get/List/Delete/Create/Patch) stay with the caller; they are out of scope
(SURVEY.md section 2.1).

The engine is anything with the two batched calls of cron_operator_b200.api.Context (UnmetSince, history_prune); the
product passes a Context (GPU), the CPU tests pass a checker built on the oracle.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from . import api


@dataclass
class CronView:
    """The fields of v1alpha1.Cron (api/v1alpha1/cron_types.go:77-107, :157) and of its owned workloads
    (cron_controller.go:136-151) that the schedule path reads.  Times are Unix seconds; None is Go's nil."""
    name: str
    schedule: str
    creation_timestamp: int
    concurrency_policy: str = "Allow"        # cron_types.go:127-139
    suspend: bool | None = None              # :96
    deadline: int | None = None              # :101
    history_limit: int | None = None         # :107
    last_schedule_time: int | None = None    # status, :157
    deletion_timestamp: int | None = None
    active: int = 0                          # len(activeWorkloads)
    finished_created_at: list = field(default_factory=list)  # creationTimestamp of every terminated workload


@dataclass
class ReconcileResult:
    """What Reconcile does for this Cron after the arithmetic (cron_controller.go:163-238, :307-346)."""
    cron: str
    requeue_after: int | None = None       # ctrl.Result{RequeueAfter: nextRun - now} (:192); None == ctrl.Result{}
    error: str | None = None               # logged by the reference, which then returns (Result{}, nil) (:185-190)
    decision: int = 0                      # api.DEC_*
    delete_active: bool = False            # Replace: delete every active workload first (:210-220)
    create: str | None = None              # name of the workload to create: "<cron>-<nextRun.Unix()>" (:222)
    set_last_schedule_time: int | None = None  # Status.LastScheduleTime = now after a create (:237)
    missed: int = 0
    too_many_missed: bool = False          # TooManyMissedTimes warning event (:431-434)
    deadline_event: bool = False           # "Deadline" event (:176-180)
    history_delete: list = field(default_factory=list)  # indices into finished_created_at to delete (:313-323)
    history_keep: list = field(default_factory=list)    # the rest, in history order (:324-344)


class BatchReconciler:
    def __init__(self, engine):
        self.engine = engine
        self._parsed = {}  # schedule text -> SpecSchedule | ParseError (the reference re-parses on every reconcile)

    def _parse(self, text):
        if text not in self._parsed:
            try:
                self._parsed[text] = api.ParseStandard(text)
            except api.ParseError as e:
                self._parsed[text] = e
        return self._parsed[text]

    def reconcile_tick(self, crons, now: int, now_nsec: int = 0):
        """One tick over the Crons that are due.  Returns one ReconcileResult per Cron, in order."""
        res = [ReconcileResult(cron=c.name) for c in crons]
        # ---- history first, for every Cron: syncStatus runs before the gates (:154-158) ----
        lens = np.array([len(c.finished_created_at) for c in crons], dtype=np.uint64)
        if lens.sum() > 0:
            seg = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
            key = np.concatenate([np.asarray(c.finished_created_at, dtype=np.int64) for c in crons if c.finished_created_at])
            lim = np.array([-1 if c.history_limit is None else c.history_limit for c in crons], dtype=np.int64)
            perm, keep = self.engine.history_prune(seg, key, lim)
            for i, r in enumerate(res):
                a, b = int(seg[i]), int(seg[i + 1])
                p, k = np.asarray(perm[a:b]), np.asarray(keep[a:b])
                r.history_delete = [int(x) for x in p[k == 0]]
                r.history_keep = [int(x) for x in p[k != 0]]
        # ---- parse; gates that need no arithmetic are left to the batched call (it applies them in Reconcile's order) --
        batch, specs = [], []
        for i, c in enumerate(crons):
            s = self._parse(c.schedule)
            if isinstance(s, api.ParseError):
                # the gates come first in Reconcile (:163-180): a deleted/suspended/expired Cron never reaches the parser
                gate = self._gate(c, now, now_nsec)
                if gate:
                    res[i].decision, res[i].deadline_event = gate, gate == api.DEC_GATE_DEADLINE
                else:
                    res[i].error = f'unparsable cron "{c.schedule}": {s}'   # :394
                    res[i].decision = api.DEC_ERR_UNSCHEDULABLE
                continue
            batch.append(i)  # either kind of schedule: "@every" rides in the same batch (api.EVERY_TAG encoding)
            specs.append(s)
        if batch:
            cs = [crons[i] for i in batch]
            earliest = [c.last_schedule_time if c.last_schedule_time is not None else c.creation_timestamp for c in cs]  # :397-402
            policy = [api.POLICY[c.concurrency_policy] for c in cs]
            active = np.array([c.active for c in cs], dtype=np.uint32)
            flags = np.array([(api.FLAG_SUSPEND if c.suspend else 0) | (api.FLAG_DELETED if c.deletion_timestamp is not None else 0)
                              for c in cs], dtype=np.uint8)
            deadline = np.array([api.ZERO if c.deadline is None else c.deadline for c in cs], dtype=np.int64)
            out = self.engine.UnmetSince(specs, earliest, now, policy, active, flags, deadline, now_nsec)
            for j, i in enumerate(batch):
                self._fan_out(res[i], crons[i], now, int(out["decision"][j]), int(out["status"][j]), int(out["last"][j]),
                              int(out["next"][j]), int(out["missed"][j]))
        return res

    @staticmethod
    def _gate(c, now, now_nsec):
        if c.deletion_timestamp is not None:
            return api.DEC_GATE_DELETED       # :163-166
        if c.suspend:
            return api.DEC_GATE_SUSPENDED     # :169-173
        if c.deadline is not None and (now > c.deadline or (now == c.deadline and now_nsec > 0)):
            return api.DEC_GATE_DEADLINE      # :176-180 (now.After(deadline))
        return 0

    @staticmethod
    def _fan_out(r, c, now, decision, status, last, nxt, missed):
        r.decision, r.missed = decision, missed
        r.too_many_missed = missed > 100
        if decision == api.DEC_GATE_DEADLINE:
            r.deadline_event = True
        if decision in (api.DEC_GATE_DELETED, api.DEC_GATE_SUSPENDED, api.DEC_GATE_DEADLINE):
            return                                              # ctrl.Result{}, nil
        if decision == api.DEC_ERR_UNSCHEDULABLE or status in (api.ST_UNSCHEDULABLE, api.ST_NONTERMINATING):
            r.error = f'unschedulable cron "{c.schedule}"'      # :411; no requeue (:185-190)
            r.too_many_missed = False
            return
        r.requeue_after = nxt - now                             # :192
        if decision in (api.DEC_CREATE, api.DEC_REPLACE_CREATE):
            r.delete_active = decision == api.DEC_REPLACE_CREATE
            r.create = api.getDefaultJobName(c.name, nxt)       # named from nextRun (:222, cron_util.go:69-71)
            r.set_last_schedule_time = now                      # :237
