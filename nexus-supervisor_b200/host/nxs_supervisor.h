// nxs_supervisor.h — host side of the classification path, in C++ (the reference is Go; no Go
// toolchain exists on this box, so the mirror of its service interface is written above the C ABI
// in C++).  Names, argument meaning and error behaviour follow services/supervisor.go of
// SneaksAndData/nexus-supervisor:
//
//   reference                                         here
//   ------------------------------------------------  -------------------------------------------
//   ProcessingConfig                   :41-47          nxs_host::ProcessingConfig
//   DecisionAction + constants         :49-56          nxs_host::ToFail... / ToRunning
//   RunStatusAnalysisResult            :58-66          nxs_host::RunStatusAnalysisResult
//   NewSupervisor(client, ns, store..) :69             Supervisor::Supervisor(...)
//   (*Supervisor).Init                 :106            Supervisor::Init
//   (*Supervisor).onEvent              :137            Supervisor::OnEvent   (flatten + enqueue)
//   (*Supervisor).superviseAction      :261            Supervisor::Flush     (GPU verdicts -> I/O)
//                                                      Supervisor::SuperviseAction (one result)
//
// What moves: the two string switches of onEvent and the decision part of superviseAction run
// on the GPU for the whole batch (nxs_classify).  What stays on the host, unchanged in meaning:
// the informer lookups (IsNexusRunEvent, GetCachedObject, labels), ReadCheckpoint before the batch,
// Jobs().Delete and UpsertCheckpoint after it.  The supervisor as a whole stays bound by those
// I/O calls (10 elements/s by default), whatever the classification costs.
#pragma once
#include <cstdint>
#include <map>
#include <optional>
#include <string>
#include <unordered_map>
#include <vector>

#include "nxs_classify.h"

namespace nxs_host {

// ---- the slices of the Kubernetes / nexus-core types the path reads
struct Event { // corev1.Event: the fields onEvent reads (:159,:171,:177-180)
  std::string involved_kind, involved_name, involved_uid, reason, message;
};
struct CachedObject { // a Job or Pod from the informer cache: only its labels matter (:181,:231-232)
  std::map<std::string, std::string> labels;
};
struct CheckpointedRequest { // nexus-core models.CheckpointedRequest: the columns the path touches
  std::string algorithm, id, lifecycle_stage, algorithm_failure_cause, algorithm_failure_details;
};

// ---- the collaborators (interfaces so tests can fake them, like the reference's fake clientset)
class ObjectCache { // resolvers.IsNexusRunEvent (:147) and resolvers.GetCachedObject[T] (:160,:211)
 public:
  virtual ~ObjectCache() = default;
  virtual bool IsNexusRunEvent(const Event &event) = 0;
  // A pointer INTO the cache (nil when the object is missing), as in the reference: read-only, not retained.
  virtual const CachedObject *GetCachedObject(const Event &event) = 0;
};
class CqlStore { // request.CqlStore: ReadCheckpoint (:264), UpsertCheckpoint (:301,:328,:353,:364)
 public:
  virtual ~CqlStore() = default;
  virtual std::optional<CheckpointedRequest> ReadCheckpoint(const std::string &algorithm, const std::string &id) = 0;
  virtual bool UpsertCheckpoint(const CheckpointedRequest &checkpoint) = 0;
};
class KubeClient { // kubeClient.BatchV1().Jobs(ns).Delete with background propagation (:268,:289,:314,:339)
 public:
  virtual ~KubeClient() = default;
  virtual bool DeleteJob(const std::string &resource_namespace, const std::string &name) = 0;
};

struct ProcessingConfig { // :41-47 (delays in milliseconds)
  int64_t FailureRateBaseDelayMs = 100;
  int64_t FailureRateMaxDelayMs = 1000;
  int RateLimitElementsPerSecond = 10;
  int RateLimitElementsBurst = 100;
  int Workers = 2;
};

using DecisionAction = std::string; // :49
extern const DecisionAction ToFailStuckInPending, ToFailFatalError, ToFailDeadlineExceeded, ToRunning; // :52-55

struct RunStatusAnalysisResult { // :58-66
  DecisionAction Action;
  std::string RunStatusMessage, RunStatusTrace, ObjectUID, ObjectKind, RequestId, Algorithm;
};

// nexus-core constants the reference only uses symbolically (values unpinned: not vendored).
struct Vocabulary {
  std::string JobTemplateNameKey = "science.sneaksanddata.com/algorithm-template-name";
  std::string JobNameLabel = "batch.kubernetes.io/job-name"; // literal in the reference (:231)
  std::string stage_names[8] = {"NEW",    "BUFFERED",          "RUNNING",           "COMPLETED",
                                "FAILED", "SCHEDULING_FAILED", "DEADLINE_EXCEEDED", "CANCELLED"};
  uint16_t finished_mask = NXS_FINISHED_MASK_DEFAULT; // stands in for (*CheckpointedRequest).IsFinished()
};

// Where the time of the last Flush went (seconds): the answer to "does the host side ever see the GPU?"
struct FlushStats {
  uint64_t events = 0, runs = 0, job_rows = 0;
  double read_checkpoints_s = 0; // ReadCheckpoint once per run (:264)
  double group_s = 0;            // records of a run made adjacent, columns filled
  double classify_s = 0;         // nxs_classify: H2D + kernel + D2H
  double apply_s = 0;            // Jobs().Delete / UpsertCheckpoint per job row
};

struct Error { // mirrors Go's `error`: empty message = nil
  std::string message;
  explicit operator bool() const { return !message.empty(); }
};

// What Flush did for one run (one job row): the observable effects of superviseAction.
struct AppliedVerdict {
  std::string algorithm, request_id;
  std::string trace; // event.Message of the winner record (-> algorithm_failure_details), "" if none
  uint32_t verdict = 0;
  bool deleted_job = false, upserted = false, skipped_finished = false;
  Error err; // "no checkpoint ..." mirrors the :265-273 path
};

class Supervisor {
 public:
  // NewSupervisor (:69): the Kubernetes client, the namespace and the store, plus the informer cache
  // and the GPU the classification runs on.
  Supervisor(KubeClient *client, std::string resource_namespace, CqlStore *cql_store, ObjectCache *cache,
             int device = 0, Vocabulary vocabulary = Vocabulary());
  ~Supervisor();
  Supervisor(const Supervisor &) = delete;
  Supervisor &operator=(const Supervisor &) = delete;

  // Init (:106): creates the GPU context.  max_batch bounds the events between two Flush calls.
  Error Init(const ProcessingConfig &config, uint64_t max_batch = 1u << 20);

  // onEvent (:137): runs the guards' lookups, interns kind/reason, resolves RequestId/Algorithm from
  // the labels and appends one record to the current batch.  Like the reference it never fails the
  // caller: events it cannot use are dropped.
  void OnEvent(const Event &event);
  size_t PendingEvents() const { return events_.size(); }
  const FlushStats &LastFlushStats() const { return stats_; }
  // Per-record verdicts are kept for AnalysisResultOf only when asked for (they cost a 4 B/record download).
  void KeepRecordVerdicts(bool keep) { keep_verdicts_ = keep; }

  // superviseAction (:261) for the whole batch: reads each run's checkpoint, classifies on the GPU,
  // then deletes jobs / upserts checkpoints as the verdicts say.  Returns a non-empty error only for
  // failures of the classification itself; per-run I/O outcomes are in `applied`.
  Error Flush(std::vector<AppliedVerdict> *applied = nullptr);

  // superviseAction (:261) for ONE analysis result, decided on the host from the same tables: the path
  // a caller takes for a single retry.  Unknown actions return the reference's error text (:372).
  Error SuperviseAction(const RunStatusAnalysisResult &result);

  // The RunStatusAnalysisResult onEvent would have sent to the actor for a record of the last batch
  // (empty Action when the event was dropped) — for tests that want to read like the reference's.
  RunStatusAnalysisResult AnalysisResultOf(size_t event_index) const;

 private:
  // One event of the batch: ids for the GPU, and where its strings sit in the string heap (SURVEY §8(a)
  // A0: msg_ref / uid_ref — the strings themselves never leave the host and are copied exactly once).
  struct Flat {
    uint32_t run;              // index into runs_
    uint32_t msg_off, msg_len; // event.Message -> RunStatusTrace
    uint32_t uid_off, uid_len; // event.InvolvedObject.UID
    uint16_t reason;
    uint8_t kind, flags;
  };
  // One run (checkpoint row) of the batch, in order of first arrival.
  struct Run {
    uint64_t job_key;
    uint32_t req_off, req_len, alg_off, alg_len; // RequestId / Algorithm in the string heap
    uint32_t count;                              // events of the run in this batch
    bool found;                                  // resolved through the informer cache
  };
  uint32_t Intern(const std::string &s);
  std::string HeapString(uint32_t off, uint32_t len) const { return heap_.substr(off, len); }
  Error Apply(const std::string &algorithm, const std::string &request_id, uint32_t verdict,
              const std::string &trace, const std::optional<CheckpointedRequest> &checkpoint, AppliedVerdict *out);
  KubeClient *client_;
  std::string ns_;
  CqlStore *store_;
  ObjectCache *cache_;
  int device_;
  Vocabulary vocab_;
  ProcessingConfig config_;
  nxs_ctx *ctx_ = nullptr;
  bool keep_verdicts_ = true;
  FlushStats stats_;
  std::string heap_, last_heap_;
  std::vector<Flat> events_, last_events_;
  std::vector<Run> runs_, last_runs_;
  std::unordered_map<uint64_t, uint32_t> run_of_;        // job key -> index into runs_
  std::unordered_map<std::string, uint8_t> stage_id_;    // lifecycle stage spelling -> id
  std::vector<uint32_t> last_verdicts_, last_pos_;       // per event of the last batch (arrival order)
  // batch columns, reused from flush to flush
  std::vector<uint8_t> col_kind_, col_flags_, col_stage_;
  std::vector<uint16_t> col_reason_;
  std::vector<uint64_t> col_key_;
  std::vector<uint32_t> col_verdict_, src_;
  std::vector<nxs_job_row> rows_;
};

uint64_t JobKey(const std::string &algorithm, const std::string &request_id); // hash of the checkpoint PK

// ---- batched checkpoint-row encoder (the step right after the classification)
// One UPDATE of nexus.checkpoints (test-resources/checkpoints.cql:1-23) per job row that calls for an
// upsert: the three columns superviseAction mutates (:297-299, :324-326, :349-351, :362), keyed by the
// row's position in the job-row array (the host maps it to (algorithm, id)).  Rows that only delete the
// job, skip or fail produce no update.  `details_record` is the batch index of the event whose Message
// becomes algorithm_failure_details, NXS_NO_WINNER when cause/details stay untouched (ToRunning).
struct CheckpointUpdate {
  uint32_t job_row;
  std::string lifecycle_stage;
  bool sets_failure = false; // false: only lifecycle_stage changes
  std::string algorithm_failure_cause;
  uint32_t details_record = NXS_NO_WINNER;
  bool delete_job = false;
};
std::vector<CheckpointUpdate> EncodeCheckpointRows(const nxs_job_row *rows, uint64_t n_rows,
                                                   const Vocabulary &vocabulary = Vocabulary());
// The algorithm_failure_cause text for a (cause_tpl, msg_tpl) pair, e.g. T1 + M2 (:325 with :198).
std::string FailureCause(uint32_t cause_tpl, uint32_t msg_tpl);

} // namespace nxs_host
