// supervisor_batch.go — the informer batching layer: how services/supervisor.go changes when the
// classification runs in batches on the GPU.  It belongs to package services of the reference
// (next to supervisor.go) and replaces two methods of *Supervisor:
//
//	onEvent          (services/supervisor.go:137-259) -> enqueueEvent + flushBatch
//	superviseAction  (services/supervisor.go:261-374) -> applyJobRow
//
// NewSupervisor, Init and Start keep their signatures (:69, :106, :376); Init additionally creates the
// gpuclassify context and starts the flush loop, the event informer registers enqueueEvent instead of
// onEvent (:124-128), and the actor's element type becomes *JobVerdict (:38, :107-117) so that the rate
// limiter and the retry back-off still guard every Delete / Upsert.
//
// NOT COMPILED HERE (no Go toolchain, un-vendored modules); gofmt-clean by inspection.  The same
// wiring is compiled and tested in C++ in nexus-supervisor_b200/host/ (nxs::Supervisor).
package services

import (
	"context"
	"fmt"
	"hash/fnv"
	"sync"
	"time"

	"github.com/SneaksAndData/nexus-core/pkg/checkpoint/models"
	"github.com/SneaksAndData/nexus-core/pkg/resolvers"
	"github.com/SneaksAndData/nexus-supervisor/services/gpuclassify"
	batchv1 "k8s.io/api/batch/v1"
	corev1 "k8s.io/api/core/v1"
	metav1 "k8s.io/apimachinery/pkg/apis/meta/v1"
	"k8s.io/apimachinery/pkg/types"
	utilruntime "k8s.io/apimachinery/pkg/util/runtime"
	"k8s.io/client-go/tools/cache"
)

// pendingEvent is one record of the next batch: the interned columns plus the strings the verdict
// needs later (they stay on the host; the GPU sees ids only).
type pendingEvent struct {
	kind      uint8
	reason    uint16
	flags     uint8
	requestId string // event.InvolvedObject.Name for Jobs (:180), the job-name label for Pods (:231)
	algorithm string // Labels[models.JobTemplateNameKey] (:181,:232)
	trace     string // event.Message -> RunStatusTrace (:177)
	uid       types.UID
	objKind   string
}

// JobVerdict is what the actor receives instead of *RunStatusAnalysisResult: one job row with the
// strings of its winner record resolved.
type JobVerdict struct {
	Row       gpuclassify.JobRow
	RequestId string
	Algorithm string
	Trace     string
	ObjectUID types.UID
}

type batcher struct {
	mu      sync.Mutex
	pending []pendingEvent
	ctx     *gpuclassify.Ctx
	batch   *gpuclassify.Batch
	// stage ids as the library numbers them (nxs_stage_name); filled once from models.LifecycleStage*
	stageID map[string]uint8
}

const (
	flagSupervised = 1 // resolvers.IsNexusRunEvent (:147)
	flagFound      = 2 // resolvers.GetCachedObject != nil (:160,:211)
	flagCheckpoint = 4 // ReadCheckpoint succeeded (:264)
	maxBatch       = 1 << 20
	flushEvery     = 50 * time.Millisecond
)

// enqueueEvent is registered as the event informer's AddFunc in place of onEvent.  It does what
// onEvent does BEFORE its switches — the sanity check, the supervised test, the cache lookups, the
// label reads — and appends one record.  It never blocks on the GPU.
func (c *Supervisor) enqueueEvent(obj interface{}) {
	if _, err := cache.ObjectToName(obj); err != nil { // :139-143
		utilruntime.HandleError(err)
		return
	}
	event := obj.(*corev1.Event)
	supervised, err := resolvers.IsNexusRunEvent(event, c.resourceNamespace, c.informers) // :147
	if err != nil {
		utilruntime.HandleError(err)
		return
	}
	rec := pendingEvent{
		kind:    gpuclassify.InternKind(event.InvolvedObject.Kind),
		reason:  gpuclassify.InternReason(event.Reason),
		trace:   event.Message,
		uid:     event.InvolvedObject.UID,
		objKind: event.InvolvedObject.Kind,
	}
	if supervised {
		rec.flags |= flagSupervised
	}
	switch event.InvolvedObject.Kind {
	case "Job": // :159-169
		job, cacheErr := resolvers.GetCachedObject[batchv1.Job](event.InvolvedObject.Name, c.resourceNamespace, c.jobInformer)
		if cacheErr != nil {
			utilruntime.HandleError(cacheErr)
			return
		}
		if job != nil {
			rec.flags |= flagFound
			rec.requestId = event.InvolvedObject.Name
			rec.algorithm = job.GetLabels()[models.JobTemplateNameKey]
		}
	case "Pod": // :210-221
		pod, cacheErr := resolvers.GetCachedObject[corev1.Pod](event.InvolvedObject.Name, c.resourceNamespace, c.podInformer)
		if cacheErr != nil {
			utilruntime.HandleError(cacheErr)
			return
		}
		if pod != nil {
			rec.flags |= flagFound
			rec.requestId = pod.GetLabels()["batch.kubernetes.io/job-name"]
			rec.algorithm = pod.GetLabels()[models.JobTemplateNameKey]
		}
	}
	c.batcher.mu.Lock()
	c.batcher.pending = append(c.batcher.pending, rec)
	full := len(c.batcher.pending) >= maxBatch
	c.batcher.mu.Unlock()
	if full {
		c.flushBatch()
	}
}

// flushLoop is started by Init next to the actor; Start's context stops it.
func (c *Supervisor) flushLoop(ctx context.Context) {
	t := time.NewTicker(flushEvery)
	defer t.Stop()
	for {
		select {
		case <-ctx.Done():
			return
		case <-t.C:
			c.flushBatch()
		}
	}
}

func runKey(algorithm, requestId string) uint64 {
	h := fnv.New64a()
	h.Write([]byte(algorithm))
	h.Write([]byte{0})
	h.Write([]byte(requestId))
	return h.Sum64()
}

// flushBatch groups the pending records by run in arrival order, reads each run's checkpoint once
// (the ReadCheckpoint of :264, moved in front of the batch), classifies the batch on the GPU and
// hands one JobVerdict per job row to the actor.
func (c *Supervisor) flushBatch() {
	b := &c.batcher
	b.mu.Lock()
	recs := b.pending
	b.pending = nil
	b.mu.Unlock()
	if len(recs) == 0 {
		return
	}
	// runs in order of first appearance, records of a run in arrival order
	order := make(map[uint64]int, len(recs))
	var runs [][]int
	keys := make([]uint64, len(recs))
	for i := range recs {
		keys[i] = runKey(recs[i].algorithm, recs[i].requestId)
		r, ok := order[keys[i]]
		if !ok {
			r = len(runs)
			order[keys[i]] = r
			runs = append(runs, nil)
		}
		runs[r] = append(runs[r], i)
	}
	src := make([]int, 0, len(recs)) // batch position -> index into recs
	n := 0
	for _, run := range runs {
		first := &recs[run[0]]
		stage, exists := uint8(0), false
		if first.flags&flagFound != 0 {
			if cp, err := c.cqlStore.ReadCheckpoint(first.algorithm, first.requestId); err == nil { // :264
				stage, exists = b.stageID[cp.LifecycleStage], true
			}
		}
		for _, i := range run {
			fl := recs[i].flags
			if exists {
				fl |= flagCheckpoint
			}
			b.batch.Kind[n], b.batch.ReasonID[n], b.batch.Flags[n] = recs[i].kind, recs[i].reason, fl
			b.batch.CurStage[n], b.batch.JobKey[n] = stage, keys[i]
			src = append(src, i)
			n++
		}
	}
	rows, err := b.ctx.Classify(b.batch, n, false)
	if err != nil { // the batch is lost like an event whose handler failed (:140-152)
		utilruntime.HandleError(err)
		return
	}
	pos := 0 // first record of the row's run in the batch
	for r, row := range rows {
		last := &recs[src[pos+len(runs[r])-1]]
		jv := &JobVerdict{Row: row, RequestId: last.requestId, Algorithm: last.algorithm, ObjectUID: last.uid}
		if row.WinnerIdx != gpuclassify.NoWinner {
			jv.Trace = recs[src[row.WinnerIdx]].trace
		}
		if row.Verdict != 0 {
			c.elementReceiverActor.Receive(jv) // rate limit, workers, retry: unchanged (:107-117)
		}
		pos += len(runs[r])
	}
}

// applyJobRow is the actor's processor in place of superviseAction: the decision has been made, what
// is left is the I/O it calls for (:265-273, :289-301, :314-328, :339-353, :362-364).
func (c *Supervisor) applyJobRow(jv *JobVerdict) (types.UID, error) {
	v := jv.Row.Verdict
	propagationPolicy := metav1.DeletePropagationBackground
	if gpuclassify.VDeleteJob(v) {
		err := c.kubeClient.BatchV1().Jobs(c.resourceNamespace).Delete(context.TODO(), jv.RequestId, metav1.DeleteOptions{
			PropagationPolicy: &propagationPolicy,
		})
		if err != nil && gpuclassify.VErr(v) == 0 {
			return jv.ObjectUID, err
		}
	}
	if gpuclassify.VErr(v) == 1 { // no checkpoint: the job is deleted, nothing is saved, the actor retries (:265-273)
		return jv.ObjectUID, fmt.Errorf("no checkpoint exists for request %s of %s", jv.RequestId, jv.Algorithm)
	}
	if !gpuclassify.VUpsert(v) { // finished run (:276-279) or nothing handled
		return jv.ObjectUID, nil
	}
	checkpoint, err := c.cqlStore.ReadCheckpoint(jv.Algorithm, jv.RequestId)
	if err != nil {
		return jv.ObjectUID, err
	}
	clone := checkpoint.DeepCopy()
	clone.LifecycleStage = c.batcher.stageName(gpuclassify.VNewStage(v))
	if tpl := gpuclassify.CauseTemplate(gpuclassify.VCauseTpl(v)); tpl != "" {
		clone.AlgorithmFailureCause = fmt.Sprintf(tpl, messageOf(jv, v)) // :298, :325, :350
		clone.AlgorithmFailureDetails = jv.Trace                         // :299, :326, :351
	}
	return jv.ObjectUID, c.cqlStore.UpsertCheckpoint(clone)
}

// messageOf returns RunStatusMessage: the three literals of the Job arms (:176,:187,:198) and
// event.Reason itself for the Pod arms (:227,:237,:247) — the template table holds both.
func messageOf(_ *JobVerdict, v uint32) string { return gpuclassify.MsgTemplate(gpuclassify.VMsgTpl(v)) }

func (b *batcher) stageName(id uint32) string {
	for name, i := range b.stageID {
		if uint32(i) == id {
			return name
		}
	}
	return ""
}
