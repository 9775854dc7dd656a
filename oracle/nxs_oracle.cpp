// nxs_oracle.cpp — CPU ORACLE.  TEST INFRASTRUCTURE ONLY.
//
// A plain C++17 restatement of the reference's failure-classification path:
//   onEvent          /root/reference/services/supervisor.go:137-259
//   superviseAction  /root/reference/services/supervisor.go:261-374
//   types            /root/reference/services/supervisor.go:49-66
// It is the checker the CUDA path is compared against and the timed CPU
// baseline.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// `--impl reference` legs may load it; the product library never does.
//
// Parity status: the two string switches and the applier switch are pinned by
// the reference SOURCE and by the 8 golden scenarios of
// services/supervisor_test.go:46-540 (tests/golden/supervisor_test_scenarios.json).
// `IsFinished()` and the LifecycleStage spellings live in the un-vendored
// dependency github.com/SneaksAndData/nexus-core v1.3.8 (go.mod:6): PARITY
// UNPINNED beyond CANCELLED=>finished, BUFFERED/RUNNING=>not finished; the
// predicate is therefore a parameter (finished_mask).  The reference itself
// cannot be built here (no Go toolchain, dependencies not vendored), so there
// is no oracle/_ref.
//
// Two modes, same outputs:
//   * string-faithful: kind/reason are strings, the switches compare strings and
//     a 7-string RunStatusAnalysisResult is built per handled event — the
//     stand-in for "the Go loop";
//   * interned: the same decision on the integer columns the GPU kernel reads.
//
// Build: see oracle/Makefile (g++ -O3 -march=native -shared -fPIC).

#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <optional>
#include <string>
#include <string_view>
#include <thread>
#include <vector>

namespace {

// ----------------------------------------------------------------- vocabulary
// Interned ids.  Reasons 1..7 are the handled ones; see the two switches.
const char *const kKindNames[3] = {"", "Job", "Pod"};
const char *const kReasonNames[16] = {
    "",               "FailedCreate", "DeadlineExceeded", "BackoffLimitExceeded",
    "PodFailurePolicy", "Started",    "Failed",           "BackOff",
    "Scheduled",      "Pulling",      "Pulled",           "Created",
    "Killing",        "SuccessfulCreate", "Completed",    "FailedScheduling"};
// Stage ids -> the spelling used by test-resources/checkpoints.cql:35,44,98 for
// the three that are pinned; the rest follow the same upper-snake convention
// (unpinned, host-side data only).
const char *const kStageNames[8] = {"NEW",    "BUFFERED",          "RUNNING",
                                    "COMPLETED", "FAILED",         "SCHEDULING_FAILED",
                                    "DEADLINE_EXCEEDED", "CANCELLED"};
enum { ST_NEW, ST_BUFFERED, ST_RUNNING, ST_COMPLETED, ST_FAILED, ST_SCHED_FAILED,
       ST_DEADLINE, ST_CANCELLED };

// DecisionAction strings, services/supervisor.go:51-56
const std::string kToFailStuckInPending = "ToFailStuckInPending";
const std::string kToFailFatalError = "ToFailFatalError";
const std::string kToFailDeadlineExceeded = "ToFailDeadlineExceeded";
const std::string kToRunning = "ToRunning";

// RunStatusMessage literals, services/supervisor.go:176,187,198
const char *const kMsgLaunch =
    "Unable to launch a container for the algorithm - please review configuration and try again.";
const char *const kMsgDeadline =
    "Algorithm exceeded its max allowed run time limit or retry attempt count.";
const char *const kMsgFatal = "Algorithm encountered a fatal error during execution.";
// AlgorithmFailureCause prefixes, services/supervisor.go:298,325
const char *const kCauseBuffered =
    "Algorithm submission was buffered, but failed to launch on the target cluster: ";
const char *const kCauseFatal = "Algorithm encountered a fatal error during execution: ";

// ------------------------------------------------------ string-faithful mode
struct Event {             // the fields of corev1.Event that onEvent reads
  std::string_view kind;   // event.InvolvedObject.Kind
  std::string_view reason; // event.Reason
  std::string_view message;
  std::string_view uid;    // event.InvolvedObject.UID
  std::string_view name;   // event.InvolvedObject.Name
};
struct CachedObject {      // what GetCachedObject + labels resolve to
  bool found;
  std::string_view job_name_label; // pod.Labels["batch.kubernetes.io/job-name"]
  std::string_view algorithm;      // Labels[models.JobTemplateNameKey]
};
struct RunStatusAnalysisResult { // services/supervisor.go:58-66
  std::string Action, RunStatusMessage, RunStatusTrace, ObjectUID, ObjectKind, RequestId,
      Algorithm;
};

// services/supervisor.go:137-259.  `supervised` = IsNexusRunEvent (:147).
std::optional<RunStatusAnalysisResult> on_event(const Event &ev, bool supervised,
                                                const CachedObject &obj) {
  if (!supervised) return std::nullopt; // :155-157
  std::optional<RunStatusAnalysisResult> out;
  if (ev.kind == "Job") { // :159
    if (!obj.found) return std::nullopt; // :161-164 stale event
    auto make = [&](const std::string &action, std::string_view msg) {
      out = RunStatusAnalysisResult{action,
                                    std::string(msg),
                                    std::string(ev.message),
                                    std::string(ev.uid),
                                    std::string(ev.kind),
                                    std::string(ev.name), // RequestId = involved name, :180
                                    std::string(obj.algorithm)};
    };
    if (ev.reason == "FailedCreate") // :172
      make(kToFailStuckInPending, kMsgLaunch);
    else if (ev.reason == "DeadlineExceeded" || ev.reason == "BackoffLimitExceeded") // :183
      make(kToFailDeadlineExceeded, kMsgDeadline);
    else if (ev.reason == "PodFailurePolicy") // :194
      make(kToFailFatalError, kMsgFatal);
    else
      return std::nullopt; // :205-206
  }
  if (ev.kind == "Pod") { // :210
    if (!obj.found) return std::nullopt; // :218-221
    auto make = [&](const std::string &action) {
      out = RunStatusAnalysisResult{action,
                                    std::string(ev.reason), // message = event.Reason, :227
                                    std::string(ev.message),
                                    std::string(ev.uid),
                                    std::string(ev.kind),
                                    std::string(obj.job_name_label), // :231
                                    std::string(obj.algorithm)};
    };
    if (ev.reason == "Started") // :224
      make(kToRunning);
    else if (ev.reason == "Failed") // :234
      make(kToFailStuckInPending);
    else if (ev.reason == "BackOff") // :244
      make(kToFailFatalError);
    // default :254-256 no-op
  }
  return out;
}

struct Checkpoint { // the fields of CheckpointedRequest the path touches
  std::string lifecycle_stage, failure_cause, failure_details;
};
struct ApplyOutcome {
  bool delete_job = false, upsert = false, skipped = false;
  int err = 0; // 1 = no checkpoint (:265-273), 2 = unknown action (:371-372)
};

// Stage ids 8..15 have no spelling in the reference; the oracle names them
// STAGE_8..STAGE_15 so the string mode covers the whole 4-bit column domain.
const char *const kExtraStageNames[8] = {"STAGE_8",  "STAGE_9",  "STAGE_10", "STAGE_11",
                                         "STAGE_12", "STAGE_13", "STAGE_14", "STAGE_15"};
const char *stage_name(uint32_t sid) {
  sid &= 15;
  return sid < 8 ? kStageNames[sid] : kExtraStageNames[sid - 8];
}
int stage_id(const std::string &s) {
  for (int i = 0; i < 16; ++i)
    if (s == stage_name(i)) return i;
  return -1;
}
// Stand-in for nexus-core's (*CheckpointedRequest).IsFinished(), :276.
bool stage_finished(const std::string &stage, uint16_t finished_mask) {
  const int sid = stage_id(stage);
  return sid >= 0 && ((finished_mask >> sid) & 1);
}

// services/supervisor.go:261-374.  `row` == nullptr models ReadCheckpoint failing.
ApplyOutcome supervise_action(const RunStatusAnalysisResult &r, Checkpoint *row,
                              uint16_t finished_mask) {
  ApplyOutcome o;
  if (row == nullptr) { // :265-273
    o.delete_job = true;
    o.err = 1;
    return o;
  }
  if (stage_finished(row->lifecycle_stage, finished_mask)) { // :276-279
    o.skipped = true;
    return o;
  }
  if (r.Action == kToFailStuckInPending) { // :284-308
    o.delete_job = true;
    row->lifecycle_stage = kStageNames[ST_SCHED_FAILED];
    row->failure_cause = std::string(kCauseBuffered) + r.RunStatusMessage;
    row->failure_details = r.RunStatusTrace;
    o.upsert = true;
  } else if (r.Action == kToFailFatalError) { // :310-335
    o.delete_job = true;
    row->lifecycle_stage = kStageNames[ST_FAILED];
    row->failure_cause = std::string(kCauseFatal) + r.RunStatusMessage;
    row->failure_details = r.RunStatusTrace;
    o.upsert = true;
  } else if (r.Action == kToFailDeadlineExceeded) { // :336-360
    o.delete_job = true;
    row->lifecycle_stage = kStageNames[ST_DEADLINE];
    row->failure_cause = r.RunStatusMessage;
    row->failure_details = r.RunStatusTrace;
    o.upsert = true;
  } else if (r.Action == kToRunning) { // :361-370
    row->lifecycle_stage = kStageNames[ST_RUNNING];
    o.upsert = true;
  } else {
    o.err = 2; // :371-372
  }
  return o;
}

// ------------------------------------------------------------ packed verdict
// Independent restatement of the bit layout documented in include/nxs_classify.h.
inline uint32_t pack(uint32_t action, uint32_t stage, uint32_t cause, uint32_t msg, uint32_t del,
                     uint32_t ups, uint32_t skip, uint32_t err) {
  return action | stage << 3 | cause << 7 | msg << 9 | del << 12 | ups << 13 | skip << 14 |
         err << 15;
}
int action_id(const std::string &a) {
  if (a == kToFailStuckInPending) return 1;
  if (a == kToFailFatalError) return 2;
  if (a == kToFailDeadlineExceeded) return 3;
  if (a == kToRunning) return 4;
  return 0;
}
int msg_tpl_id(const std::string &m) {
  if (m == kMsgLaunch) return 1;
  if (m == kMsgDeadline) return 2;
  if (m == kMsgFatal) return 3;
  if (m == "Started") return 4;
  if (m == "Failed") return 5;
  if (m == "BackOff") return 6;
  return 0;
}
int cause_tpl_id(const std::string &cause, const std::string &msg) {
  if (cause == std::string(kCauseBuffered) + msg) return 1;
  if (cause == std::string(kCauseFatal) + msg) return 2;
  if (cause == msg) return 3;
  return 0;
}

struct JobRow {
  uint64_t job_key;
  uint32_t verdict;
  uint32_t winner_idx;
};
struct Columns {
  const uint8_t *kind;
  const uint16_t *reason;
  const uint8_t *flags;
  const uint8_t *stage;
  const uint64_t *key;
};

// State of one job while its records are replayed in order.
struct JobFold {
  uint32_t last_action = 0, stage = 0, cause = 0, msg = 0, del = 0, ups = 0, skip = 0, err = 0;
  uint32_t winner = 0xFFFFFFFFu;
  uint32_t row() const { return pack(last_action, ups ? stage : 0, cause, msg, del, ups, skip, err); }
};

// String-faithful replay of records [lo, hi) (hi ends on a job boundary or n).
// Per record: rebuild the strings the Go code sees, run on_event, then apply the
// result twice — once against a row holding the record's own cur_stage (the
// per-record verdict) and once against the job's evolving row (the fold).
uint64_t classify_strings(const Columns &c, uint64_t lo, uint64_t hi, uint64_t idx_base,
                          uint16_t finished_mask, uint32_t *verdict, JobRow *rows) {
  uint64_t n_rows = 0;
  JobFold fold;
  Checkpoint job_row;
  char name_buf[40], uid_buf[40];
  for (uint64_t i = lo; i < hi; ++i) {
    const bool head = (i == lo) || c.key[i] != c.key[i - 1];
    if (head) {
      fold = JobFold();
      job_row = Checkpoint{stage_name(c.stage[i]), "", ""};
    }
    const uint8_t kind = c.kind[i];
    const uint16_t reason = c.reason[i];
    const uint8_t flags = c.flags[i];
    snprintf(name_buf, sizeof name_buf, "%016llx", (unsigned long long)c.key[i]);
    snprintf(uid_buf, sizeof uid_buf, "uid-%llu", (unsigned long long)i);
    Event ev{kind < 3 ? kKindNames[kind] : "Other", reason < 16 ? kReasonNames[reason] : "Unknown",
             "", uid_buf, name_buf};
    CachedObject obj{(flags & 2) != 0, name_buf, "test-algorithm"};
    auto res = on_event(ev, (flags & 1) != 0, obj);
    uint32_t v = 0;
    if (res) {
      const uint32_t a = action_id(res->Action), m = msg_tpl_id(res->RunStatusMessage);
      const bool exists = (flags & 4) != 0;
      // (1) per-record verdict: apply against a row holding the record's own cur_stage
      Checkpoint own{stage_name(c.stage[i]), "", ""};
      const ApplyOutcome o = supervise_action(*res, exists ? &own : nullptr, finished_mask);
      v = pack(a, o.upsert ? stage_id(own.lifecycle_stage) : 0,
               o.upsert && a != 4 ? cause_tpl_id(own.failure_cause, res->RunStatusMessage) : 0, m,
               o.delete_job, o.upsert, o.skipped, o.err);
      // (2) the job fold: apply against the job's evolving row
      const ApplyOutcome f = supervise_action(*res, exists ? &job_row : nullptr, finished_mask);
      fold.del |= f.delete_job; fold.ups |= f.upsert; fold.skip |= f.skipped;
      if (f.err) fold.err = f.err;
      if (f.upsert) {
        fold.last_action = a;
        fold.stage = stage_id(job_row.lifecycle_stage);
        if (a != 4) { // ToRunning leaves cause/details untouched (:361-364)
          fold.cause = cause_tpl_id(job_row.failure_cause, res->RunStatusMessage);
          fold.msg = m;
          fold.winner = (uint32_t)(idx_base + i);
        }
      }
    }
    if (verdict) verdict[i] = v;
    const bool tail = (i + 1 == hi) || c.key[i + 1] != c.key[i];
    if (tail && rows) rows[n_rows] = JobRow{c.key[i], fold.row(), fold.winner};
    n_rows += tail;
  }
  return n_rows;
}

// ------------------------------------------------------------- interned mode
// The same decision on ids.  Table rows = the switch arms of onEvent.
struct Arm { uint8_t kind, reason, action, msg; };
const Arm kArms[7] = {
    {1, 1, 1, 1}, // Job FailedCreate         -> StuckInPending, M0   :172-182
    {1, 2, 3, 2}, // Job DeadlineExceeded     -> DeadlineExceeded, M1 :183-193
    {1, 3, 3, 2}, // Job BackoffLimitExceeded -> DeadlineExceeded, M1 :183-193
    {1, 4, 2, 3}, // Job PodFailurePolicy     -> FatalError, M2       :194-204
    {2, 5, 4, 4}, // Pod Started              -> Running              :224-233
    {2, 6, 1, 5}, // Pod Failed               -> StuckInPending       :234-243
    {2, 7, 2, 6}, // Pod BackOff              -> FatalError           :244-253
};
// applier switch, services/supervisor.go:283-370: action -> (stage, cause template, delete)
const uint8_t kNewStage[5] = {0, ST_SCHED_FAILED, ST_FAILED, ST_DEADLINE, ST_RUNNING};
const uint8_t kCauseTpl[5] = {0, 1, 2, 3, 0};

inline void lookup(uint8_t kind, uint16_t reason, uint8_t flags, uint32_t &action, uint32_t &msg) {
  action = msg = 0;
  if ((flags & 3) != 3) return; // !supervised (:155) or object missing (:161,:218)
  for (const Arm &a : kArms)
    if (a.kind == kind && a.reason == reason) { action = a.action; msg = a.msg; return; }
}

uint64_t classify_interned(const Columns &c, uint64_t lo, uint64_t hi, uint64_t idx_base,
                           uint16_t finished_mask, uint32_t *verdict, JobRow *rows) {
  uint64_t n_rows = 0;
  JobFold fold;
  uint32_t job_stage = 0;
  for (uint64_t i = lo; i < hi; ++i) {
    const bool head = (i == lo) || c.key[i] != c.key[i - 1];
    if (head) { fold = JobFold(); job_stage = c.stage[i] & 15; }
    uint32_t a, m;
    lookup(c.kind[i], c.reason[i], c.flags[i], a, m);
    uint32_t v = 0;
    if (a) {
      const bool exists = (c.flags[i] & 4) != 0;
      const uint32_t own = c.stage[i] & 15;
      if (!exists) v = pack(a, 0, 0, m, 1, 0, 0, 1);
      else if ((finished_mask >> own) & 1) v = pack(a, 0, 0, m, 0, 0, 1, 0);
      else v = pack(a, kNewStage[a], kCauseTpl[a], m, a != 4, 1, 0, 0);
      if (!exists) { fold.err = 1; fold.del = 1; }
      else if ((finished_mask >> job_stage) & 1) fold.skip = 1;
      else {
        job_stage = kNewStage[a];
        fold.ups = 1; fold.last_action = a; fold.stage = job_stage;
        if (a != 4) { fold.del = 1; fold.cause = kCauseTpl[a]; fold.msg = m;
                      fold.winner = (uint32_t)(idx_base + i); }
      }
    }
    if (verdict) verdict[i] = v;
    const bool tail = (i + 1 == hi) || c.key[i + 1] != c.key[i];
    if (tail && rows) rows[n_rows] = JobRow{c.key[i], fold.row(), fold.winner};
    n_rows += tail;
  }
  return n_rows;
}

// ----------------------------------------------------- synthetic generator
// Counter-based stream; the CUDA generator implements the same spec (DESIGN.md
// "Synthetic workloads").
inline uint64_t splitmix64(uint64_t x) {
  uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
struct Rec { uint8_t kind; uint16_t reason; uint8_t flags; uint8_t stage; uint64_t key; };

const uint8_t kPairKind[8] = {1, 1, 1, 1, 2, 2, 2, 2};
const uint8_t kPairReason[8] = {1, 2, 3, 4, 5, 6, 7, 8};

Rec gen_uniform(uint64_t seed, uint64_t i) {
  const uint64_t r = splitmix64(seed ^ i);
  Rec o;
  o.kind = kPairKind[r & 7];
  o.reason = kPairReason[r & 7];
  o.flags = 7;
  o.stage = ((r >> 8) & 1) ? ST_RUNNING : ST_BUFFERED;
  o.key = splitmix64(r);
  return o;
}

const uint32_t kJobPeriod = 61;      // a forced job boundary every 61 records bounds look-back
const uint32_t kHeadThreshold = 26214; // 0.4 * 65536: P(new job starts at a record)
inline bool raw_head(uint64_t seed, uint64_t i) {
  return (i % kJobPeriod) == 0 || (splitmix64(seed ^ i) & 0xFFFF) < kHeadThreshold;
}
inline uint64_t job_head(uint64_t seed, uint64_t i) {
  uint64_t j = i;
  while (!raw_head(seed, j)) --j;
  return j + ((i - j) & ~7ull); // at most 8 records per job
}
// failure tail, Zipf(s=1.2) over 6 (kind, reason) pairs: cumulative 16-bit thresholds
const uint32_t kZipfCum[6] = {30427, 43672, 51813, 57580, 61991, 65536};
const uint8_t kZipfKind[6] = {2, 2, 1, 1, 1, 1};
const uint8_t kZipfReason[6] = {7, 6, 4, 3, 2, 1};
const uint8_t kFinishedStages[5] = {ST_COMPLETED, ST_FAILED, ST_SCHED_FAILED, ST_DEADLINE,
                                    ST_CANCELLED};

Rec gen_skewed(uint64_t seed, uint64_t i) {
  const uint64_t c = splitmix64(seed ^ i);
  const uint64_t h = job_head(seed, i);
  const uint64_t key = splitmix64(splitmix64(seed ^ h) ^ 0xA5A5A5A5A5A5A5A5ull);
  Rec o;
  o.key = key;
  const uint32_t u = (c >> 16) & 0xFFFF, u2 = (c >> 32) & 0xFFFF, u3 = (uint32_t)(c >> 48);
  if (u < 58982) {          // 90 % healthy
    if (u2 < 21845) { o.kind = 2; o.reason = 5; } // Pod/Started (30 % of all records)
    else {
      o.reason = 8 + (u2 % 7);                    // unhandled reasons 8..14
      o.kind = o.reason >= 13 ? 1 : 2;
    }
  } else {
    int k = 0;
    while (u3 >= kZipfCum[k]) ++k;
    o.kind = kZipfKind[k];
    o.reason = kZipfReason[k];
  }
  // per job: 2 % already finished, else BUFFERED/RUNNING; 0.1 % no checkpoint
  const uint32_t sv = (key >> 8) & 0xFFFF;
  if (sv < 1311) o.stage = kFinishedStages[(key >> 24) % 5];
  else o.stage = ((key >> 32) & 1) ? ST_RUNNING : ST_BUFFERED;
  uint8_t flags = 7;
  if (((key >> 40) & 0x3FF) == 0) flags &= ~4;
  if (((c >> 40) & 0x3FF) == 0) flags &= ~2; // 0.1 % stale events, per record
  o.flags = flags;
  return o;
}

} // namespace

// =================================================================== C ABI
// ------------------------------------------------- timed stand-in for the Go loop
// What services/supervisor.go does per event, and nothing else: the strings of the event and of the
// cached object EXIST already when onEvent runs (they are fields of *corev1.Event and of the
// informer's objects, :145,:160,:211), so they are built here before the clock starts.  Timed: the
// two string switches and the 7-string RunStatusAnalysisResult per handled event (onEvent, :159-257),
// then superviseAction's decision against the run's checkpoint row held in memory (:264-370:
// IsFinished, the action switch, the Sprintf of the cause) — one row per job, replayed in record
// order.  The three network calls of superviseAction are not modelled.
struct EventStrings {
  std::string name, uid; // event.InvolvedObject.Name / UID
};
struct LoopResult {
  uint64_t handled = 0, upserts = 0, deletes = 0, checksum = 0;
};
LoopResult event_loop(const Columns &c, const EventStrings *strs, uint64_t lo, uint64_t hi,
                      uint16_t finished_mask) {
  LoopResult r;
  Checkpoint job_row;
  static const std::string_view kAlgorithm = "test-algorithm";
  for (uint64_t i = lo; i < hi; ++i) {
    if (i == lo || c.key[i] != c.key[i - 1]) job_row = Checkpoint{stage_name(c.stage[i]), "", ""};
    const uint8_t kind = c.kind[i], flags = c.flags[i];
    const uint16_t reason = c.reason[i];
    const EventStrings &es = strs[i - lo];
    const Event ev{kind < 3 ? kKindNames[kind] : "Other", reason < 16 ? kReasonNames[reason] : "Unknown", "",
                   es.uid, es.name};
    const CachedObject obj{(flags & 2) != 0, es.name, kAlgorithm};
    const auto res = on_event(ev, (flags & 1) != 0, obj);
    if (!res) continue;
    ++r.handled;
    const ApplyOutcome o = supervise_action(*res, (flags & 4) ? &job_row : nullptr, finished_mask);
    r.upserts += o.upsert;
    r.deletes += o.delete_job;
    r.checksum += job_row.failure_cause.size() + res->RequestId.size();
  }
  return r;
}

extern "C" {

// String-mode replay that returns what the reference would have WRITTEN: for every job whose row was
// upserted at least once, the three strings superviseAction set (lifecycle_stage,
// algorithm_failure_cause, algorithm_failure_details), back to back in `arena`, offsets[3u..3u+3] per
// updated job u, update_job[u] = index of the job in the stream.  event.Message of record i is the text
// "m<i>".  Returns the number of updated jobs (UINT64_MAX when the arena is too small).
uint64_t nxo_encode_strings(const uint8_t *kind, const uint16_t *reason, const uint8_t *flags, const uint8_t *stage,
                            const uint64_t *key, uint64_t n, uint16_t finished_mask, uint8_t *arena, uint64_t cap,
                            uint64_t *offsets, uint32_t *update_job, uint64_t *arena_bytes) {
  uint64_t n_up = 0, pos = 0, job = 0;
  Checkpoint row;
  bool upserted = false;
  char name_buf[40], msg_buf[32];
  for (uint64_t i = 0; i < n; ++i) {
    if (i == 0 || key[i] != key[i - 1]) {
      row = Checkpoint{stage_name(stage[i]), "", ""};
      upserted = false;
    }
    snprintf(name_buf, sizeof name_buf, "%016llx", (unsigned long long)key[i]);
    snprintf(msg_buf, sizeof msg_buf, "m%llu", (unsigned long long)i);
    const Event ev{kind[i] < 3 ? kKindNames[kind[i]] : "Other", reason[i] < 16 ? kReasonNames[reason[i]] : "Unknown",
                   msg_buf, "uid", name_buf};
    const CachedObject obj{(flags[i] & 2) != 0, name_buf, "test-algorithm"};
    const auto res = on_event(ev, (flags[i] & 1) != 0, obj);
    if (res) upserted |= supervise_action(*res, (flags[i] & 4) ? &row : nullptr, finished_mask).upsert;
    if (i + 1 == n || key[i + 1] != key[i]) {
      if (upserted) {
        const std::string *parts[3] = {&row.lifecycle_stage, &row.failure_cause, &row.failure_details};
        for (int k = 0; k < 3; ++k) {
          if (pos + parts[k]->size() > cap) return UINT64_MAX;
          offsets[3 * n_up + k] = pos;
          memcpy(arena + pos, parts[k]->data(), parts[k]->size());
          pos += parts[k]->size();
        }
        update_job[n_up++] = (uint32_t)job;
      }
      ++job;
    }
  }
  offsets[3 * n_up] = pos;
  *arena_bytes = pos;
  return n_up;
}

// Best-of-`repeats` wall time of event_loop over the n records, split over `threads` ranges cut on job
// boundaries; the per-record strings are materialised by each thread before the timed region.
double nxo_bench_event_loop(int threads, const uint8_t *kind, const uint16_t *reason, const uint8_t *flags,
                            const uint8_t *stage, const uint64_t *key, uint64_t n, uint16_t finished_mask,
                            int repeats, uint64_t *handled, uint64_t *checksum) {
  if (threads < 1) threads = 1;
  Columns c{kind, reason, flags, stage, key};
  std::vector<uint64_t> cut(threads + 1, n);
  cut[0] = 0;
  for (int t = 1; t < threads; ++t) {
    uint64_t p = n / threads * t;
    if (p < cut[t - 1]) p = cut[t - 1];
    while (p < n && p > 0 && key[p] == key[p - 1]) ++p;
    cut[t] = p;
  }
  std::vector<std::vector<EventStrings>> strs(threads);
  std::vector<LoopResult> results(threads);
  std::atomic<int> ready{0}, go{0}, done{0};
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&, t] {
      const uint64_t lo = cut[t], hi = cut[t + 1];
      strs[t].resize(hi - lo);
      char buf[40];
      for (uint64_t i = lo; i < hi; ++i) { // (untimed) what the informer cache already holds
        snprintf(buf, sizeof buf, "%016llx", (unsigned long long)key[i]);
        strs[t][i - lo].name = buf;
        snprintf(buf, sizeof buf, "uid-%llu", (unsigned long long)i);
        strs[t][i - lo].uid = buf;
      }
      for (int rep = 1; rep <= repeats; ++rep) {
        ready.fetch_add(1);
        while (go.load(std::memory_order_acquire) < rep) std::this_thread::yield();
        results[t] = event_loop(c, strs[t].data(), lo, hi, finished_mask);
        done.fetch_add(1);
      }
    });
  double best = 1e30;
  for (int rep = 1; rep <= repeats; ++rep) {
    while (ready.load() < threads * rep) std::this_thread::yield();
    const auto t0 = std::chrono::steady_clock::now();
    go.store(rep, std::memory_order_release);
    while (done.load() < threads * rep) std::this_thread::yield();
    const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (dt < best) best = dt;
  }
  for (auto &th : pool) th.join();
  uint64_t h = 0, cs = 0;
  for (const LoopResult &r : results) h += r.handled, cs += r.checksum + r.upserts + r.deletes;
  if (handled) *handled = h;
  if (checksum) *checksum = cs;
  return best;
}

} // extern "C"

extern "C" {

struct nxo_job_row { uint64_t job_key; uint32_t verdict; uint32_t winner_idx; };

// mode 0 = interned, 1 = string-faithful.  threads <= 1 runs inline.  With
// threads > 1 the records are split into ranges that end on job boundaries and
// the per-range rows are concatenated in order.
uint64_t nxo_classify(int mode, int threads, const uint8_t *kind, const uint16_t *reason,
                      const uint8_t *flags, const uint8_t *stage, const uint64_t *key, uint64_t n,
                      uint16_t finished_mask, uint32_t *verdict, nxo_job_row *rows) {
  Columns c{kind, reason, flags, stage, key};
  auto run = mode == 1 ? classify_strings : classify_interned;
  static_assert(sizeof(JobRow) == sizeof(nxo_job_row), "row layout");
  JobRow *out_rows = reinterpret_cast<JobRow *>(rows);
  if (threads <= 1 || n < 4096) return run(c, 0, n, 0, finished_mask, verdict, out_rows);
  std::vector<uint64_t> cut(threads + 1, n);
  cut[0] = 0;
  for (int t = 1; t < threads; ++t) {
    uint64_t p = n / threads * t;
    if (p < cut[t - 1]) p = cut[t - 1];
    while (p < n && p > 0 && key[p] == key[p - 1]) ++p;
    cut[t] = p;
  }
  std::vector<uint64_t> counts(threads, 0);
  std::vector<std::vector<JobRow>> parts(threads);
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&, t] {
      const uint64_t lo = cut[t], hi = cut[t + 1];
      if (hi <= lo) return;
      if (out_rows) parts[t].resize(hi - lo);
      // Range-local replay: shift every column so the range starts at 0.
      Columns s{kind + lo, reason + lo, flags + lo, stage + lo, key + lo};
      counts[t] = run(s, 0, hi - lo, lo, finished_mask, verdict ? verdict + lo : nullptr,
                      out_rows ? parts[t].data() : nullptr);
    });
  for (auto &th : pool) th.join();
  uint64_t total = 0;
  for (int t = 0; t < threads; ++t) {
    if (out_rows && counts[t]) memcpy(out_rows + total, parts[t].data(), counts[t] * sizeof(JobRow));
    total += counts[t];
  }
  return total;
}

// Synthetic stream records [first, first+n); keeps only `shard` of `n_shards`
// (splitmix64(job_key) % n_shards) when n_shards > 1.  Returns records written.
uint64_t nxo_generate(int workload, uint64_t seed, uint64_t first, uint64_t n, uint32_t shard,
                      uint32_t n_shards, uint8_t *kind, uint16_t *reason, uint8_t *flags,
                      uint8_t *stage, uint64_t *key, uint64_t capacity) {
  uint64_t w = 0;
  for (uint64_t i = first; i < first + n; ++i) {
    const Rec r = workload == 1 ? gen_uniform(seed, i) : gen_skewed(seed, i);
    if (n_shards > 1 && splitmix64(r.key) % n_shards != shard) continue;
    if (w >= capacity) return ~0ull;
    kind[w] = r.kind; reason[w] = r.reason; flags[w] = r.flags; stage[w] = r.stage; key[w] = r.key;
    ++w;
  }
  return w;
}

uint64_t nxo_splitmix64(uint64_t x) { return splitmix64(x); }

// One event through the string API, for the golden scenarios: returns the final
// lifecycle stage / cause / details in caller buffers (each `cap` bytes).
// found_ckpt = 0 models a failing ReadCheckpoint.  Return: bit0 delete_job,
// bit1 upsert, bit2 skipped, bits 4-5 err, bit 8 = a result was produced.
int nxo_event_strings(const char *kind, const char *reason, const char *message, int supervised,
                      int found, int found_ckpt, const char *stage_in, uint16_t finished_mask,
                      char *stage_out, char *cause_out, char *details_out, char *action_out,
                      char *msg_out, size_t cap) {
  Event ev{kind, reason, message, "uid", "request-id"};
  CachedObject obj{found != 0, "request-id", "test-algorithm"};
  auto res = on_event(ev, supervised != 0, obj);
  Checkpoint row{stage_in, "", ""};
  snprintf(stage_out, cap, "%s", stage_in);
  cause_out[0] = details_out[0] = action_out[0] = msg_out[0] = 0;
  if (!res) return 0;
  ApplyOutcome o = supervise_action(*res, found_ckpt ? &row : nullptr, finished_mask);
  snprintf(stage_out, cap, "%s", row.lifecycle_stage.c_str());
  snprintf(cause_out, cap, "%s", row.failure_cause.c_str());
  snprintf(details_out, cap, "%s", row.failure_details.c_str());
  snprintf(action_out, cap, "%s", res->Action.c_str());
  snprintf(msg_out, cap, "%s", res->RunStatusMessage.c_str());
  return (o.delete_job ? 1 : 0) | (o.upsert ? 2 : 0) | (o.skipped ? 4 : 0) | (o.err << 4) | 256;
}

// supervise_action on an arbitrary action string (the `default:` arm, :371-372).
int nxo_apply_action_string(const char *action, const char *stage_in, uint16_t finished_mask) {
  RunStatusAnalysisResult r{action, "m", "t", "u", "Job", "id", "alg"};
  Checkpoint row{stage_in, "", ""};
  return supervise_action(r, &row, finished_mask).err;
}

const char *nxo_kind_name(uint32_t i) { return i < 3 ? kKindNames[i] : ""; }
const char *nxo_reason_name(uint32_t i) { return i < 16 ? kReasonNames[i] : ""; }
const char *nxo_stage_name(uint32_t i) { return i < 8 ? kStageNames[i] : ""; }
unsigned nxo_hardware_threads(void) { return std::thread::hardware_concurrency(); }

} // extern "C"
