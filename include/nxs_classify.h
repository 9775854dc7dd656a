/*
 * nxs_classify.h — C ABI of the B200 failure-classification library.
 *
 * This is the drop-in boundary for ONE path of SneaksAndData/nexus-supervisor:
 * the event classifier `(*Supervisor).onEvent` (services/supervisor.go:137-259)
 * followed by the decision part of `(*Supervisor).superviseAction`
 * (services/supervisor.go:261-374), applied to a batch of event records that the
 * host has flattened into struct-of-arrays columns.  The reference exposes no
 * FFI of its own (it is pure Go, CGO_ENABLED=0); the entry points below are what
 * a cgo shim next to services/supervisor.go would bind.
 *
 * Conventions
 *   - every function returns 0 (NXS_OK) or a negative nxs_status; nothing aborts
 *     or throws across the boundary; nxs_strerror() names a status;
 *   - plain pointers and sizes only; the caller owns every buffer and the library
 *     reads/writes it only during the call (for *_device calls: until the ctx
 *     stream reaches the end of the work the call enqueued);
 *   - a ctx belongs to one GPU and one caller thread at a time;
 *   - there is no CPU fallback: without a CUDA device nxs_create fails.
 *
 * All citations `services/...:N` are relative to the reference repository.
 */
#ifndef NXS_CLASSIFY_H
#define NXS_CLASSIFY_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NXS_ABI_VERSION 1

/* ---------------------------------------------------------------- status */
typedef enum nxs_status {
  NXS_OK = 0,
  NXS_ERR_ARG = -1,       /* null pointer, misaligned column, bad enum       */
  NXS_ERR_CAPACITY = -2,  /* n exceeds the ctx capacity / 2^32-2 records     */
  NXS_ERR_CUDA = -3,      /* a CUDA runtime call failed (see nxs_last_error) */
  NXS_ERR_NO_DEVICE = -4, /* no usable CUDA device: there is no CPU fallback */
  NXS_ERR_NOMEM = -5,
  NXS_ERR_UNKNOWN_ACTION = -6, /* services/supervisor.go:371-372             */
  NXS_ERR_NOALLOC_LIMIT = -7   /* n > NXS_NOALLOC_MAX_RECORDS (static scratch) */
} nxs_status;

/* ------------------------------------------------- interned vocabularies */
/* event.InvolvedObject.Kind (services/supervisor.go:159,210) */
enum { NXS_KIND_OTHER = 0, NXS_KIND_JOB = 1, NXS_KIND_POD = 2 };

/* event.Reason, interned.  1..7 are the seven reasons the two switches handle
 * (services/supervisor.go:171-207 and :223-257); the rest are common kubelet /
 * job-controller reasons that fall into `default`.  0 and every id >= 32 are
 * "unknown" and also fall into `default`. */
enum {
  NXS_REASON_UNKNOWN = 0,
  NXS_REASON_FAILED_CREATE = 1,          /* Job, :172 */
  NXS_REASON_DEADLINE_EXCEEDED = 2,      /* Job, :183 */
  NXS_REASON_BACKOFF_LIMIT_EXCEEDED = 3, /* Job, :183 */
  NXS_REASON_POD_FAILURE_POLICY = 4,     /* Job, :194 */
  NXS_REASON_STARTED = 5,                /* Pod, :224 */
  NXS_REASON_FAILED = 6,                 /* Pod, :234 */
  NXS_REASON_BACKOFF = 7,                /* Pod, :244 */
  NXS_REASON_SCHEDULED = 8,
  NXS_REASON_PULLING = 9,
  NXS_REASON_PULLED = 10,
  NXS_REASON_CREATED = 11,
  NXS_REASON_KILLING = 12,
  NXS_REASON_SUCCESSFUL_CREATE = 13,
  NXS_REASON_COMPLETED = 14,
  NXS_REASON_FAILED_SCHEDULING = 15,
  NXS_REASON_LUT_SIZE = 32
};

/* checkpoint.LifecycleStage as an id (strings live in nexus-core, host-side) */
enum {
  NXS_STAGE_NEW = 0,
  NXS_STAGE_BUFFERED = 1,
  NXS_STAGE_RUNNING = 2,
  NXS_STAGE_COMPLETED = 3,
  NXS_STAGE_FAILED = 4,
  NXS_STAGE_SCHEDULING_FAILED = 5,
  NXS_STAGE_DEADLINE_EXCEEDED = 6,
  NXS_STAGE_CANCELLED = 7
};

/* Default for the unpinned nexus-core predicate `IsFinished()`
 * (services/supervisor.go:276): bit s set <=> stage s counts as finished.
 * Pinned by the reference test: CANCELLED finished, BUFFERED/RUNNING not. */
#define NXS_FINISHED_MASK_DEFAULT                                            \
  ((uint16_t)((1u << NXS_STAGE_COMPLETED) | (1u << NXS_STAGE_FAILED) |       \
              (1u << NXS_STAGE_SCHEDULING_FAILED) |                          \
              (1u << NXS_STAGE_DEADLINE_EXCEEDED) | (1u << NXS_STAGE_CANCELLED)))

/* DecisionAction (services/supervisor.go:49-56) */
enum {
  NXS_ACTION_NONE = 0,
  NXS_ACTION_STUCK_IN_PENDING = 1,  /* ToFailStuckInPending   */
  NXS_ACTION_FATAL_ERROR = 2,       /* ToFailFatalError       */
  NXS_ACTION_DEADLINE_EXCEEDED = 3, /* ToFailDeadlineExceeded */
  NXS_ACTION_RUNNING = 4            /* ToRunning              */
};

/* flags column */
#define NXS_FLAG_SUPERVISED 0x01u /* resolvers.IsNexusRunEvent, :147-157      */
#define NXS_FLAG_FOUND 0x02u      /* resolvers.GetCachedObject != nil, :160/211 */
#define NXS_FLAG_CHECKPOINT 0x04u /* ReadCheckpoint succeeded, :264-273       */

/* ---------------------------------------------------------------- verdict
 * One u32 per record and one per job row.
 * The PER-RECORD verdict is what the record would get on its own against the checkpoint stage
 * in its own cur_stage: it ignores the earlier records of the same job, so records that follow
 * a stop of their job still carry upsert / delete_job.  It is diagnostic.  I/O (Jobs().Delete,
 * UpsertCheckpoint) must be driven from the JOB ROWS only: they hold the outcome of the
 * reference's sequential replay with its IsFinished() guard (services/supervisor.go:276-279).
 *   bits  0-2  action        DecisionAction id
 *   bits  3-6  new_stage     stage id written by the upsert (0 unless upsert)
 *   bits  7-8  cause_tpl     nxs_cause_template index (0 = cause untouched)
 *   bits  9-11 msg_tpl       nxs_msg_template index (RunStatusMessage)
 *   bit   12   delete_job    Jobs(ns).Delete issued (:268,:289,:314,:339)
 *   bit   13   upsert        UpsertCheckpoint issued (:301,:328,:353,:364)
 *   bit   14   skip_finished IsFinished() no-op (:276-279)
 *   bits 15-16 err           1 = no checkpoint (:265-273), 2 = unknown action
 */
#define NXS_V_ACTION(v) ((uint32_t)(v) & 7u)
#define NXS_V_NEW_STAGE(v) (((uint32_t)(v) >> 3) & 15u)
#define NXS_V_CAUSE_TPL(v) (((uint32_t)(v) >> 7) & 3u)
#define NXS_V_MSG_TPL(v) (((uint32_t)(v) >> 9) & 7u)
#define NXS_V_DELETE_JOB(v) (((uint32_t)(v) >> 12) & 1u)
#define NXS_V_UPSERT(v) (((uint32_t)(v) >> 13) & 1u)
#define NXS_V_SKIP_FINISHED(v) (((uint32_t)(v) >> 14) & 1u)
#define NXS_V_ERR(v) (((uint32_t)(v) >> 15) & 3u)
#define NXS_V_PACK(action, stage, cause, msg, del, ups, skip, err)           \
  ((uint32_t)(action) | ((uint32_t)(stage) << 3) | ((uint32_t)(cause) << 7) | \
   ((uint32_t)(msg) << 9) | ((uint32_t)(del) << 12) | ((uint32_t)(ups) << 13) | \
   ((uint32_t)(skip) << 14) | ((uint32_t)(err) << 15))
#define NXS_ERR_NO_CHECKPOINT 1u
#define NXS_NO_WINNER 0xFFFFFFFFu

/* Per-job row: the outcome of replaying, in record order, every record of one
 * job against its checkpoint row — what the reference does one event at a time
 * through ReadCheckpoint / IsFinished / switch / UpsertCheckpoint
 * (services/supervisor.go:264-370).  Records of a job must be adjacent.
 *   verdict.action     action of the LAST applied record (decides the stage)
 *   verdict.new_stage  resulting lifecycle stage (if upsert)
 *   verdict.cause_tpl / msg_tpl  from the winner = the last applied record
 *                      with a failure action; its message/trace become
 *                      algorithm_failure_cause / _details (:298-299,:325-326,:350-351)
 *   verdict.delete_job / upsert / skip_finished / err   OR over the records
 *   winner_idx         record index of the winner in this batch, NXS_NO_WINNER if none
 */
typedef struct nxs_job_row {
  uint64_t job_key;
  uint32_t verdict;
  uint32_t winner_idx;
} nxs_job_row;

/* Input columns, one element per event record. */
typedef struct nxs_batch {
  const uint8_t *kind;       /* NXS_KIND_*                                   */
  const uint16_t *reason_id; /* NXS_REASON_*                                 */
  const uint8_t *flags;      /* NXS_FLAG_*                                   */
  const uint8_t *cur_stage;  /* NXS_STAGE_* of the row ReadCheckpoint gave   */
  const uint64_t *job_key;   /* hash of (algorithm, id) — the checkpoint PK  */
  uint64_t n;                /* records; < 2^32 - 1                          */
} nxs_batch;

/* The same record in 10 bytes instead of 13, for callers whose batches cross PCIe: the four small
 * columns folded into one u16.  Interning already bounds the values, so nothing is lost:
 *   bits  0-5  reason id, ids >= 63 stored as 63   (ids >= 32 all fall into `default`, :205,:254)
 *   bits  6-7  kind, kinds >= 3 stored as 3        (neither Job nor Pod, :159,:210)
 *   bits  8-10 flags (NXS_FLAG_*)
 *   bits 11-14 checkpoint stage id & 15
 * The verdict of a packed record is the verdict of the record it was packed from. */
#define NXS_ATTR_PACK(kind, reason, flags, stage)                                                   \
  (((uint32_t)(reason) > 63u ? 63u : (uint32_t)(reason)) | ((uint32_t)(kind) > 3u ? 3u : (uint32_t)(kind)) << 6 | \
   ((uint32_t)(flags) & 7u) << 8 | ((uint32_t)(stage) & 15u) << 11)
#define NXS_ATTR_REASON(a) ((uint32_t)(a) & 63u)
#define NXS_ATTR_KIND(a) (((uint32_t)(a) >> 6) & 3u)
#define NXS_ATTR_FLAGS(a) (((uint32_t)(a) >> 8) & 7u)
#define NXS_ATTR_STAGE(a) (((uint32_t)(a) >> 11) & 15u)
typedef struct nxs_batch_packed {
  const uint64_t *job_key; /* as in nxs_batch                                */
  const uint16_t *attr;    /* NXS_ATTR_PACK(kind, reason_id, flags, cur_stage) */
  uint64_t n;
} nxs_batch_packed;

typedef struct nxs_out {
  uint32_t *verdict;     /* n; may be NULL: per-record verdicts are then not produced
                            (every entry point that takes an nxs_out)                 */
  nxs_job_row *job_rows; /* capacity >= number of jobs (n is always enough) */
  uint64_t *n_jobs;      /* 1 element                                       */
} nxs_out;

typedef struct nxs_ctx nxs_ctx;

/* ------------------------------------------------------------ life cycle */
/* Replaces the per-Supervisor state built in NewSupervisor/Init
 * (services/supervisor.go:69-135): one ctx per (GPU, caller thread).
 * `finished_mask` stands in for nexus-core's IsFinished() (:276). */
int nxs_create(int device, uint64_t max_records, uint16_t finished_mask,
               nxs_ctx **out);
void nxs_destroy(nxs_ctx *ctx);
/* Use an existing cudaStream_t (passed as void*) instead of the ctx's own.  Work already queued
 * by this ctx on its previous stream is waited for first (launches of one ctx share scratch).
 * Every entry point leaves the calling thread's current CUDA device as it found it. */
int nxs_set_stream(nxs_ctx *ctx, void *cuda_stream);
int nxs_sync(nxs_ctx *ctx);
const char *nxs_strerror(int status);
const char *nxs_last_error(const nxs_ctx *ctx);
int nxs_abi_version(void);

/* ------------------------------------------------------------- hot path */
/* onEvent + superviseAction decision for a batch (services/supervisor.go:137-374).
 * Host pointers; copies in, classifies on the GPU, copies verdicts, job rows and
 * the job count back before returning.  Uses a few host threads of its own for large batches
 * (copies through page-locked staging when the caller's memory is pageable; for batches of more
 * than one 8 Mi-record slice, packing of the small columns before the upload and expansion of
 * the verdicts, which then cross the link as one byte each, after the download). */
int nxs_classify(nxs_ctx *ctx, const nxs_batch *in, const nxs_out *out);
/* The same for a packed batch (host pointers): 10 instead of 13 bytes per record go up, and with
 * out->verdict == NULL only the job rows come down — the layout for callers behind PCIe, who act
 * on job rows anyway.  The attr column is expanded into the four byte columns on the device. */
int nxs_classify_packed(nxs_ctx *ctx, const nxs_batch_packed *in, const nxs_out *out);
/* Bytes the last nxs_classify / nxs_classify_packed call of this ctx moved over the host link, as queued
 * (measurement aid: batches of more than one slice upload the four small columns packed into 16 bits —
 * 10 instead of 13 bytes per record, the kernel reads no more of them — and download 1 instead of 4 bytes
 * per verdict). */
void nxs_last_wire_bytes(const nxs_ctx *ctx, uint64_t *h2d_bytes, uint64_t *d2h_bytes);
/* Host helper of the flattening layer: attr[i] = NXS_ATTR_PACK(kind[i], reason_id[i], flags[i], cur_stage[i]). */
void nxs_pack_columns(const uint8_t *kind, const uint16_t *reason_id, const uint8_t *flags,
                      const uint8_t *cur_stage, uint64_t n, uint16_t *attr);
/* Same with DEVICE pointers (16-byte aligned columns); asynchronous on the ctx
 * stream.  out->n_jobs is a device pointer too. */
int nxs_classify_device(nxs_ctx *ctx, const nxs_batch *d_in,
                        const nxs_out *d_out);
/* Number of kernels the ctx has launched so far (bench.py's gpu_launches). */
uint64_t nxs_launch_count(const nxs_ctx *ctx);
/* Records per tile of the kernel variant the last classification launch of the ctx used
 * (the library picks one of its built-in tile sizes from the batch size); 0 before any. */
int nxs_last_tile(const nxs_ctx *ctx);

/* No counterpart in the reference.  The same classification as
 * nxs_classify_device, for hosts that must not let a
 * library allocate device memory — e.g. a service that gave its whole HBM budget
 * to its own pool at start-up, so a cudaMalloc from here would fail or eat into
 * it, or a latency-critical loop that cannot afford the device-wide stall a
 * cudaMalloc/cudaFree can cause.  It needs no nxs_ctx and calls no allocating
 * routine: the tile descriptors and counters the kernel uses as scratch live in
 * a static __device__ array inside the library (about 32 KiB on every GPU that
 * loads the library, until the process exits), which bounds a call to
 * NXS_NOALLOC_MAX_RECORDS records; larger batches are refused with
 * NXS_ERR_NOALLOC_LIMIT and nothing is launched.
 *   - d_in / d_out hold DEVICE pointers with the same alignment rules as
 *     nxs_classify_device; the work is asynchronous on `cuda_stream`.
 *   - The first call on a device resolves the scratch array's device address and
 *     configures the kernel for that device; make that first call before any
 *     stream capture begins.
 *   - All calls on a device share the one scratch array (each call clears the
 *     part it uses first), so two calls whose GPU work can overlap — different
 *     streams, or different threads — race; serialise them. */
#define NXS_NOALLOC_MAX_RECORDS (1u << 24)
int nxs_classify_device_noalloc(int device, void *cuda_stream, uint16_t finished_mask,
                                const nxs_batch *d_in, const nxs_out *d_out);

/* pinned host memory for nxs_classify callers */
int nxs_alloc_host(size_t bytes, void **out);
void nxs_free_host(void *p);

/* --------------------------------------------------- host-side vocabulary */
/* Interning used by the host flattening layer: what the Go side would do with
 * event.InvolvedObject.Kind / event.Reason before filling the columns. */
uint8_t nxs_intern_kind(const char *s, size_t len);
uint16_t nxs_intern_reason(const char *s, size_t len);
const char *nxs_reason_name(uint32_t reason_id);
const char *nxs_action_name(uint32_t action); /* :52-55 */
const char *nxs_stage_name(uint32_t stage);
/* RunStatusMessage texts (:176,:187,:198,:227,:237,:247); idx 0 = "" */
const char *nxs_msg_template(uint32_t idx);
/* AlgorithmFailureCause formats (:298,:325,:350); idx 0 = NULL (untouched) */
const char *nxs_cause_template(uint32_t idx);

/* EXTENSION — no counterpart in the reference, no parity claimed.  The reference classifies Kubernetes
 * Events; container-status reasons (OOMKilled, ImagePullBackOff, ...) never reach it as such.  For
 * streams phrased in that vocabulary this names the EVENT through which kubelet / the job controller
 * surface such a status (outside the reference): OOMKilled -> Job / PodFailurePolicy (exit 137 matching
 * the Job's podFailurePolicy: the reference's own comment, services/supervisor.go:311-312),
 * ImagePullBackOff and CrashLoopBackOff -> Pod / BackOff, ErrImagePull -> Pod / Failed.  Returns 1 and
 * fills *kind / *reason_id for those four, 0 (and kind = NXS_KIND_OTHER, reason NXS_REASON_UNKNOWN)
 * for anything else.  Exit codes, restart counts and phases have no semantics in the reference and
 * are not inputs anywhere in this library. */
int nxs_ext_container_reason_to_event(const char *container_reason, size_t len, uint8_t *kind, uint16_t *reason_id);

/* ------------------------------------------- checkpoint-row encoder (host) */
/* The step right after the classification: for every job row that calls for an upsert, the three
 * columns of nexus.checkpoints (test-resources/checkpoints.cql:1-23) that superviseAction sets
 * (services/supervisor.go:297-299, :324-326, :349-351, :362) as bytes, back to back in one arena:
 *   update u (u = 0 .. *n_updates-1, in job-row order) belongs to job row update_row[u];
 *   lifecycle_stage          = arena[offsets[3u]   .. offsets[3u+1])   stage name of verdict.new_stage
 *   algorithm_failure_cause  = arena[offsets[3u+1] .. offsets[3u+2])   cause format applied to the message text
 *   algorithm_failure_details= arena[offsets[3u+2] .. offsets[3u+3])   Message of the winner event
 *   update_flags[u] (optional) bit 0: cause / details are SET (else the row keeps its own: ToRunning,
 *   :361-364, and both ranges are empty); bit 1: the job is also deleted.
 * Rows that only delete the job, skip or fail produce no update.  msg_heap / msg_offsets (n_records + 1
 * entries) hold event.Message per batch record; NULL = details empty.  stage_names: 16 spellings, or
 * NULL for nxs_stage_name().  With arena == NULL and offsets == NULL only *n_updates and *arena_bytes
 * are computed (size query); offsets needs 3 * n_updates + 1 entries.  Host memory, multi-threaded. */
int nxs_encode_checkpoint_rows(const nxs_job_row *rows, uint64_t n_rows, const uint8_t *msg_heap,
                               const uint64_t *msg_offsets, uint64_t n_records, const char *const *stage_names,
                               uint8_t *arena, uint64_t arena_capacity, uint64_t *offsets, uint32_t *update_row,
                               uint8_t *update_flags, uint64_t *n_updates, uint64_t *arena_bytes);

/* ------------------------------------------------------------- sharding */
/* Shard of a job for a G-GPU run: splitmix64(job_key) % n_shards. */
uint32_t nxs_shard_of(uint64_t job_key, uint32_t n_shards);

/* One host batch on several GPUs.  The unit of dependence of the path is one job = one
 * checkpoint row — superviseAction reads and writes only the row of its (algorithm, id)
 * (services/supervisor.go:264,:301; primary key test-resources/checkpoints.cql:22) — so a batch
 * can be split by job: nxs_classify_sharded partitions the records by nxs_shard_of(job_key,
 * n_shards) without ever splitting a job, classifies every shard with its own ctx through
 * nxs_classify (shard s runs on devices[s % n_devices]: one GPU per shard when there are
 * enough, round-robin otherwise), and puts the results back in stream order on the host:
 * out->verdict[i] belongs to record i, out->job_rows are ordered by the position at which
 * each job ends in the batch and their winner_idx are batch indices — bit for bit what one
 * nxs_classify call over the whole batch returns.  No collective and no peer access is
 * involved: the "gather" is G independent device-to-host copies plus a host merge.
 * Host pointers; the handle owns pinned staging buffers that grow to the largest shard seen.
 * Not thread-safe per handle. */
typedef struct nxs_sharded nxs_sharded;
int nxs_sharded_create(uint32_t n_shards, const int *devices, uint32_t n_devices, uint64_t max_records,
                       uint16_t finished_mask, nxs_sharded **out);
void nxs_sharded_destroy(nxs_sharded *h);
int nxs_classify_sharded(nxs_sharded *h, const nxs_batch *in, const nxs_out *out);
const char *nxs_sharded_last_error(const nxs_sharded *h);

/* ------------------------------------------------- synthetic workloads */
enum { NXS_WORKLOAD_UNIFORM = 1, NXS_WORKLOAD_SKEWED = 2 };
#define NXS_DEFAULT_SEED 0x6E65787573ull
/* Writable view of the input columns (device pointers). */
typedef struct nxs_columns {
  uint8_t *kind;
  uint16_t *reason_id;
  uint8_t *flags;
  uint8_t *cur_stage;
  uint64_t *job_key;
  uint64_t capacity;
} nxs_columns;
/* Fill device columns with records [first, first+n) of the counter-based
 * synthetic stream (bench / test input; the oracle regenerates the same stream
 * on the CPU).  With n_shards > 1 only records whose job falls in `shard` are
 * kept, compacted in stream order; *n_kept (host pointer) receives the count. */
int nxs_generate_device(nxs_ctx *ctx, int workload, uint64_t seed,
                        uint64_t first, uint64_t n, uint32_t shard,
                        uint32_t n_shards, const nxs_columns *d_cols,
                        uint64_t *n_kept);

#ifdef __cplusplus
}
#endif
#endif /* NXS_CLASSIFY_H */
