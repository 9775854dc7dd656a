// Package gpuclassify binds libnxs_classify.so (include/nxs_classify.h) into nexus-supervisor.
//
// NOT COMPILED HERE: neither this repository's build container nor its GPU boxes have a Go
// toolchain, and the reference's modules (nexus-core v1.3.8, k8s.io/*) are not vendored.  The file is
// kept gofmt-clean by inspection and is written to be correct under cgo's pointer-passing rules.  The
// same C ABI is exercised from Python (nexus-supervisor_b200/_abi.py) and from C++
// (nexus-supervisor_b200/host/), which is what the tests of this repository run.
//
// The reference has no FFI for this path (pure Go, CGO_ENABLED=0, .container/Dockerfile:14); this
// package is what a maintainer would add as services/gpuclassify, together with CGO_ENABLED=1 and the
// shared object in the image.
//
// cgo rule that shapes the code: a Go pointer passed to C must not point to memory that itself holds
// Go pointers, and C must not keep Go pointers after the call.  nxs_batch / nxs_out are structs OF
// pointers, so they and every column they point to live in C memory here: the structs come from
// C.calloc, the columns from nxs_alloc_host (page-locked, which is also what makes the host-to-device
// copies run at PCIe speed).  Go code fills the columns through slices that alias that C memory
// (unsafe.Slice); no Go-allocated memory ever crosses the boundary.
package gpuclassify

/*
#cgo CFLAGS: -I${SRCDIR}/../../third_party/nxs/include
#cgo LDFLAGS: -L${SRCDIR}/../../third_party/nxs/lib -lnxs_classify
#include <stdlib.h>
#include "nxs_classify.h"
*/
import "C"

import (
	"fmt"
	"unsafe"
)

// Verdict field accessors (bit layout: include/nxs_classify.h, NXS_V_*).
func VAction(v uint32) uint32 { return v & 7 }
func VNewStage(v uint32) uint32 { return (v >> 3) & 15 }
func VCauseTpl(v uint32) uint32 { return (v >> 7) & 3 }
func VMsgTpl(v uint32) uint32 { return (v >> 9) & 7 }
func VDeleteJob(v uint32) bool { return (v>>12)&1 != 0 }
func VUpsert(v uint32) bool { return (v>>13)&1 != 0 }
func VSkipFinished(v uint32) bool { return (v>>14)&1 != 0 }
func VErr(v uint32) uint32 { return (v >> 15) & 3 }

const NoWinner = 0xFFFFFFFF

// JobRow mirrors nxs_job_row (16 bytes, same field order).
type JobRow struct {
	JobKey    uint64
	Verdict   uint32
	WinnerIdx uint32
}

// Ctx wraps one nxs_ctx: one GPU, used from one goroutine at a time (services/supervisor.go builds
// its per-Supervisor state in NewSupervisor/Init, :69-135; this is the GPU part of it).
type Ctx struct{ p *C.nxs_ctx }

func New(device int, maxRecords uint64, finishedMask uint16) (*Ctx, error) {
	var p *C.nxs_ctx // a Go variable holding a C pointer: passing &p is allowed (it holds no Go pointer)
	if rc := C.nxs_create(C.int(device), C.uint64_t(maxRecords), C.uint16_t(finishedMask), &p); rc != C.NXS_OK {
		return nil, fmt.Errorf("nxs_create: %s", C.GoString(C.nxs_strerror(rc)))
	}
	return &Ctx{p}, nil
}

func (c *Ctx) Close() { C.nxs_destroy(c.p); c.p = nil }

// Batch owns the flattened columns of up to Cap records, the outputs and the two argument structs —
// all in C memory.  The exported slices alias that memory; fill them in place, then call Classify.
type Batch struct {
	Cap      int
	Kind     []uint8  // InternKind(event.InvolvedObject.Kind)            (:159,:210)
	ReasonID []uint16 // InternReason(event.Reason)                       (:171,:223)
	Flags    []uint8  // supervised | found<<1 | checkpointExists<<2      (:147,:160,:211,:264)
	CurStage []uint8  // stage id of the checkpoint read for the run      (:264)
	JobKey   []uint64 // hash of (algorithm, requestId), runs adjacent
	Verdict  []uint32 // per-record verdicts (diagnostic)
	Rows     []JobRow // one row per run: what drives Delete / Upsert

	in    *C.nxs_batch
	out   *C.nxs_out
	nJobs *C.uint64_t
	bufs  []unsafe.Pointer
}

func pinned(bytes int) (unsafe.Pointer, error) {
	var p unsafe.Pointer
	if rc := C.nxs_alloc_host(C.size_t(bytes), &p); rc != C.NXS_OK {
		return nil, fmt.Errorf("nxs_alloc_host(%d): %s", bytes, C.GoString(C.nxs_strerror(rc)))
	}
	return p, nil
}

func NewBatch(capacity int) (*Batch, error) {
	b := &Batch{Cap: capacity}
	sizes := []int{capacity, 2 * capacity, capacity, capacity, 8 * capacity, 4 * capacity, 16 * capacity}
	for _, s := range sizes {
		p, err := pinned(s)
		if err != nil {
			b.Free()
			return nil, err
		}
		b.bufs = append(b.bufs, p)
	}
	b.Kind = unsafe.Slice((*uint8)(b.bufs[0]), capacity)
	b.ReasonID = unsafe.Slice((*uint16)(b.bufs[1]), capacity)
	b.Flags = unsafe.Slice((*uint8)(b.bufs[2]), capacity)
	b.CurStage = unsafe.Slice((*uint8)(b.bufs[3]), capacity)
	b.JobKey = unsafe.Slice((*uint64)(b.bufs[4]), capacity)
	b.Verdict = unsafe.Slice((*uint32)(b.bufs[5]), capacity)
	b.Rows = unsafe.Slice((*JobRow)(b.bufs[6]), capacity)
	b.in = (*C.nxs_batch)(C.calloc(1, C.size_t(unsafe.Sizeof(C.nxs_batch{}))))
	b.out = (*C.nxs_out)(C.calloc(1, C.size_t(unsafe.Sizeof(C.nxs_out{}))))
	b.nJobs = (*C.uint64_t)(C.calloc(1, 8))
	b.in.kind = (*C.uint8_t)(b.bufs[0])
	b.in.reason_id = (*C.uint16_t)(b.bufs[1])
	b.in.flags = (*C.uint8_t)(b.bufs[2])
	b.in.cur_stage = (*C.uint8_t)(b.bufs[3])
	b.in.job_key = (*C.uint64_t)(b.bufs[4])
	b.out.verdict = (*C.uint32_t)(b.bufs[5])
	b.out.job_rows = (*C.nxs_job_row)(b.bufs[6])
	b.out.n_jobs = b.nJobs
	return b, nil
}

func (b *Batch) Free() {
	for _, p := range b.bufs {
		C.nxs_free_host(p)
	}
	b.bufs = nil
	C.free(unsafe.Pointer(b.in))
	C.free(unsafe.Pointer(b.out))
	C.free(unsafe.Pointer(b.nJobs))
	b.in, b.out, b.nJobs = nil, nil, nil
}

// Classify runs onEvent's switches and the decision part of superviseAction for records [0, n) of the
// batch (services/supervisor.go:137-374) and returns the job rows, which alias b.Rows.  Every pointer
// handed to C points to C memory, so no pinning is needed and cgocheck has nothing to object to.
func (c *Ctx) Classify(b *Batch, n int, wantVerdicts bool) ([]JobRow, error) {
	if n < 0 || n > b.Cap {
		return nil, fmt.Errorf("gpuclassify: n=%d outside the batch capacity %d", n, b.Cap)
	}
	b.in.n = C.uint64_t(n)
	if wantVerdicts {
		b.out.verdict = (*C.uint32_t)(b.bufs[5])
	} else {
		b.out.verdict = nil // rows only: the per-record verdicts are neither produced nor copied back
	}
	if rc := C.nxs_classify(c.p, b.in, b.out); rc != C.NXS_OK {
		return nil, fmt.Errorf("nxs_classify: %s: %s", C.GoString(C.nxs_strerror(rc)), C.GoString(C.nxs_last_error(c.p)))
	}
	return b.Rows[:int(*b.nJobs)], nil
}

// Interning: what enqueueEvent does with event.InvolvedObject.Kind / event.Reason before filling the
// columns.  The string's bytes are only read during the call; unsafe.StringData of a Go string is a
// pointer to pointer-free memory, which cgo allows.
func InternKind(s string) uint8 {
	return uint8(C.nxs_intern_kind((*C.char)(unsafe.Pointer(unsafe.StringData(s))), C.size_t(len(s))))
}

func InternReason(s string) uint16 {
	return uint16(C.nxs_intern_reason((*C.char)(unsafe.Pointer(unsafe.StringData(s))), C.size_t(len(s))))
}

func MsgTemplate(i uint32) string { return C.GoString(C.nxs_msg_template(C.uint32_t(i))) }
func StageName(i uint32) string { return C.GoString(C.nxs_stage_name(C.uint32_t(i))) }
func ActionName(i uint32) string { return C.GoString(C.nxs_action_name(C.uint32_t(i))) }
func ShardOf(key uint64, g uint32) uint32 { return uint32(C.nxs_shard_of(C.uint64_t(key), C.uint32_t(g))) }

// CauseTemplate returns the fmt verb string of AlgorithmFailureCause (:298,:325,:350); "" = untouched.
func CauseTemplate(i uint32) string {
	p := C.nxs_cause_template(C.uint32_t(i))
	if p == nil {
		return ""
	}
	return C.GoString(p)
}
