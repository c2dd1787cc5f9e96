// reconciler.go -- a batching front for internal/controller's Reconcile (SURVEY.md section 8f rank 2), the Go form of
// cron_operator_b200/reconciler.py (which is the tested one: tests/test_reconciler.py).
//
// UNVERIFIED SOURCE, like the rest of this package: never compiled (no Go toolchain on the build image).
//
// The reference reconciles one Cron per call (cron_controller.go:90-238).  Everything in that call that is
// arithmetic -- parse, the sched.Next() chain from the last schedule time to now, the concurrency decision, the
// history order and prune -- is the same for every Cron, so a tick takes the Crons the work queue has ready, makes one
// UnmetSince and one HistoryPrune call, and hands back per Cron what Reconcile does next.  The Kubernetes calls stay
// with the caller.
package b200cron

import (
	"fmt"
	"time"
)

// CronView carries the fields of v1alpha1.Cron (api/v1alpha1/cron_types.go:77-107, :157) and of its owned
// workloads (cron_controller.go:136-151) that the schedule path reads.
type CronView struct {
	Name              string
	Schedule          string
	CreationTimestamp time.Time
	ConcurrencyPolicy string // "Allow", "Forbid", "Replace" (cron_types.go:127-139)
	Suspend           bool
	Deadline          *time.Time
	HistoryLimit      *int64
	LastScheduleTime  *time.Time
	Deleted           bool        // DeletionTimestamp != nil
	Active            int         // len(activeWorkloads)
	FinishedCreatedAt []time.Time // creationTimestamp of every terminated workload
}

// TickResult says what Reconcile does for one Cron after the arithmetic (cron_controller.go:163-238, :307-346).
type TickResult struct {
	Cron                string
	Requeue             bool          // ctrl.Result{RequeueAfter: ...} (:192); false == ctrl.Result{}
	RequeueAfter        time.Duration
	Err                 error         // logged by the reference, which then returns (Result{}, nil) (:185-190)
	Decision            uint8         // B200CRON_DEC_*
	DeleteActive        bool          // Replace: delete every active workload first (:210-220)
	Create              string        // workload name "<cron>-<nextRun.Unix()>" (:222); "" == nothing to create
	SetLastScheduleTime bool          // Status.LastScheduleTime = now after a create (:237)
	Missed              uint32
	TooManyMissed       bool          // TooManyMissedTimes warning event (:431-434)
	DeadlineEvent       bool          // "Deadline" event (:176-180)
	HistoryDelete       []int         // indices into FinishedCreatedAt to delete (:313-323)
	HistoryKeep         []int         // the rest, in history order (:324-344)
}

const (
	decGateDeleted = 1 + iota
	decGateSuspended
	decGateDeadline
	decErrUnschedulable
	decRequeueOnly
	decSkipForbid
	decReplaceCreate
	decCreate
)

var policyCode = map[string]uint8{"": 0, "Allow": 0, "Forbid": 1, "Replace": 2}

// BatchReconciler caches parsed schedules (the reference re-parses on every reconcile, :392).
type BatchReconciler struct {
	Engine *Engine
	parsed map[string]*Schedule
	bad    map[string]error
}

func NewBatchReconciler(e *Engine) *BatchReconciler {
	return &BatchReconciler{Engine: e, parsed: map[string]*Schedule{}, bad: map[string]error{}}
}

func (r *BatchReconciler) parse(text string) (*Schedule, error) {
	if s, ok := r.parsed[text]; ok {
		return s, nil
	}
	if err, ok := r.bad[text]; ok {
		return nil, err
	}
	s, err := r.Engine.ParseStandard(text)
	if err != nil {
		r.bad[text] = err
		return nil, err
	}
	r.parsed[text] = s
	return s, nil
}

func gate(c *CronView, now time.Time) uint8 {
	switch {
	case c.Deleted:
		return decGateDeleted // :163-166
	case c.Suspend:
		return decGateSuspended // :169-173
	case c.Deadline != nil && now.After(*c.Deadline):
		return decGateDeadline // :176-180
	}
	return 0
}

// ReconcileTick evaluates every Cron of one tick with one UnmetSince and one HistoryPrune call.
func (r *BatchReconciler) ReconcileTick(crons []CronView, now time.Time) ([]TickResult, error) {
	res := make([]TickResult, len(crons))
	for i := range crons {
		res[i].Cron = crons[i].Name
	}
	// history first, for every Cron: syncStatus runs before the gates (:154-158)
	segOff := make([]uint64, len(crons)+1)
	var created, limit []int64
	for i := range crons {
		for _, t := range crons[i].FinishedCreatedAt {
			created = append(created, t.Unix())
		}
		segOff[i+1] = uint64(len(created))
		if crons[i].HistoryLimit != nil {
			limit = append(limit, *crons[i].HistoryLimit)
		} else {
			limit = append(limit, -1)
		}
	}
	if len(created) > 0 {
		perm, keep, err := r.Engine.HistoryPrune(segOff, created, limit)
		if err != nil {
			return nil, err
		}
		for i := range crons {
			for j := segOff[i]; j < segOff[i+1]; j++ {
				if keep[j] == 0 {
					res[i].HistoryDelete = append(res[i].HistoryDelete, int(perm[j]))
				} else {
					res[i].HistoryKeep = append(res[i].HistoryKeep, int(perm[j]))
				}
			}
		}
	}
	var batch []int
	var specs []Spec
	var earliest, deadline []int64
	var policy, flags []uint8
	var active []uint32
	for i := range crons {
		c := &crons[i]
		s, err := r.parse(c.Schedule)
		if err != nil || s.Every > 0 {
			// the gates come first in Reconcile: a deleted, suspended or expired Cron never reaches the parser
			if g := gate(c, now); g != 0 {
				res[i].Decision, res[i].DeadlineEvent = g, g == decGateDeadline
				continue
			}
			if err != nil {
				res[i].Err = fmt.Errorf("unparsable cron %q: %w", c.Schedule, err) // :394
				res[i].Decision = decErrUnschedulable
				continue
			}
			r.every(&res[i], c, s.Every, now)
			continue
		}
		batch = append(batch, i)
		specs = append(specs, s.Spec)
		e := c.CreationTimestamp // :397-402
		if c.LastScheduleTime != nil {
			e = *c.LastScheduleTime
		}
		earliest = append(earliest, e.Unix())
		policy = append(policy, policyCode[c.ConcurrencyPolicy])
		active = append(active, uint32(c.Active))
		var f uint8
		if c.Suspend {
			f |= 1
		}
		if c.Deleted {
			f |= 2
		}
		flags = append(flags, f)
		if c.Deadline != nil {
			deadline = append(deadline, c.Deadline.Unix())
		} else {
			deadline = append(deadline, ZeroTime)
		}
	}
	if len(batch) > 0 {
		out, err := r.Engine.UnmetSince(specs, earliest, now, policy, active, flags, deadline)
		if err != nil {
			return nil, err
		}
		for j, i := range batch {
			fanOut(&res[i], &crons[i], now, out[j])
		}
	}
	return res, nil
}

// every answers a ConstantDelaySchedule on the host (constantdelay.go: Next(t) = t + delay on whole seconds).
func (r *BatchReconciler) every(out *TickResult, c *CronView, delay int64, now time.Time) {
	e := c.CreationTimestamp
	if c.LastScheduleTime != nil {
		e = *c.LastScheduleTime
	}
	res := Result{LastMissed: ZeroTime, Next: now.Unix() + delay}
	if !e.After(now) {
		if missed := (now.Unix() - e.Unix()) / delay; missed > 0 {
			res.Missed = uint32(missed)
			res.LastMissed = e.Unix() + missed*delay
		}
	}
	switch {
	case res.LastMissed == ZeroTime:
		res.Decision = decRequeueOnly
	case policyCode[c.ConcurrencyPolicy] == 1 && c.Active > 0:
		res.Decision = decSkipForbid
	case policyCode[c.ConcurrencyPolicy] == 2:
		res.Decision = decReplaceCreate
	default:
		res.Decision = decCreate
	}
	fanOut(out, c, now, res)
}

func fanOut(out *TickResult, c *CronView, now time.Time, r Result) {
	out.Decision, out.Missed = r.Decision, r.Missed
	out.TooManyMissed = r.Missed > 100
	switch r.Decision {
	case decGateDeadline:
		out.DeadlineEvent = true
		return
	case decGateDeleted, decGateSuspended:
		return // ctrl.Result{}, nil
	}
	if r.Decision == decErrUnschedulable || r.Status == 1 || r.Status == 3 {
		out.Err = fmt.Errorf("unschedulable cron %q", c.Schedule) // :411; no requeue (:185-190)
		out.TooManyMissed = false
		return
	}
	out.Requeue = true
	out.RequeueAfter = time.Unix(r.Next, 0).Sub(now) // :192
	if r.Decision == decCreate || r.Decision == decReplaceCreate {
		out.DeleteActive = r.Decision == decReplaceCreate
		out.Create = fmt.Sprintf("%s-%d", c.Name, r.Next) // named from nextRun (:222, cron_util.go:69-71)
		out.SetLastScheduleTime = true                    // :237
	}
}
