// nxs_supervisor.cpp — see nxs_supervisor.h.  Citations: services/supervisor.go of the reference.
#include "nxs_supervisor.h"

#include <algorithm>
#include <chrono>
#include <cstring>
#include <cstdio>
#include <numeric>
#include <unordered_map>

namespace nxs_host {

const DecisionAction ToFailStuckInPending = "ToFailStuckInPending";     // :52
const DecisionAction ToFailFatalError = "ToFailFatalError";             // :53
const DecisionAction ToFailDeadlineExceeded = "ToFailDeadlineExceeded"; // :54
const DecisionAction ToRunning = "ToRunning";                           // :55

uint64_t JobKey(const std::string &algorithm, const std::string &request_id) {
  uint64_t h = 0xcbf29ce484222325ull; // FNV-1a over "algorithm\0id", then a splitmix64 finaliser
  auto eat = [&](const std::string &s) {
    for (unsigned char c : s) h = (h ^ c) * 0x100000001b3ull;
    h = (h ^ 0xffu) * 0x100000001b3ull;
  };
  eat(algorithm);
  eat(request_id);
  h += 0x9E3779B97F4A7C15ull;
  h = (h ^ (h >> 30)) * 0xBF58476D1CE4E5B9ull;
  h = (h ^ (h >> 27)) * 0x94D049BB133111EBull;
  return h ^ (h >> 31);
}

Supervisor::Supervisor(KubeClient *client, std::string resource_namespace, CqlStore *cql_store, ObjectCache *cache,
                       int device, Vocabulary vocabulary)
    : client_(client), ns_(std::move(resource_namespace)), store_(cql_store), cache_(cache), device_(device),
      vocab_(std::move(vocabulary)) {}

Supervisor::~Supervisor() { nxs_destroy(ctx_); }

Error Supervisor::Init(const ProcessingConfig &config, uint64_t max_batch) {
  config_ = config;
  nxs_destroy(ctx_);
  ctx_ = nullptr;
  stage_id_.clear();
  for (uint8_t s = 0; s < 8; ++s) stage_id_[vocab_.stage_names[s]] = s;
  const int rc = nxs_create(device_, max_batch, vocab_.finished_mask, &ctx_);
  if (rc != NXS_OK) return Error{std::string("nxs_create: ") + nxs_strerror(rc)};
  return Error{};
}

uint32_t Supervisor::Intern(const std::string &s) {
  const uint32_t off = (uint32_t)heap_.size();
  heap_.append(s);
  return off;
}

namespace {
double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
} // namespace

// onEvent, :137-259, minus the two switches (they run on the GPU).
void Supervisor::OnEvent(const Event &event) {
  Flat f{};
  const bool supervised = cache_->IsNexusRunEvent(event); // :147
  if (supervised) f.flags |= NXS_FLAG_SUPERVISED;
  f.kind = nxs_intern_kind(event.involved_kind.data(), event.involved_kind.size());
  f.reason = nxs_intern_reason(event.reason.data(), event.reason.size());
  const std::string *request_id = nullptr, *algorithm = nullptr;
  static const std::string kEmpty;
  if (supervised && (f.kind == NXS_KIND_JOB || f.kind == NXS_KIND_POD)) {
    if (const CachedObject *obj = cache_->GetCachedObject(event)) { // :160, :211
      f.flags |= NXS_FLAG_FOUND;
      auto label = [&](const std::string &key) -> const std::string * {
        const auto it = obj->labels.find(key);
        return it == obj->labels.end() ? &kEmpty : &it->second;
      };
      algorithm = label(vocab_.JobTemplateNameKey);                                              // :181, :232
      request_id = f.kind == NXS_KIND_JOB ? &event.involved_name : label(vocab_.JobNameLabel); // :180, :231
    }
  }
  // Events that cannot lead to a verdict still travel (the kernel drops them): they keep their place
  // in the batch, and their key groups them with nothing that has a checkpoint.
  static const std::string kUnresolved = "\x01unresolved";
  const uint64_t key = request_id ? JobKey(*algorithm, *request_id) : JobKey(kUnresolved, event.involved_kind + "/" + event.involved_name);
  auto it = run_of_.find(key);
  if (it == run_of_.end()) {
    Run r{};
    r.job_key = key;
    r.found = request_id != nullptr;
    if (request_id) {
      r.req_off = Intern(*request_id), r.req_len = (uint32_t)request_id->size();
      r.alg_off = Intern(*algorithm), r.alg_len = (uint32_t)algorithm->size();
    }
    it = run_of_.emplace(key, (uint32_t)runs_.size()).first;
    runs_.push_back(r);
  }
  f.run = it->second;
  runs_[f.run].count += 1;
  f.msg_off = Intern(event.message), f.msg_len = (uint32_t)event.message.size();
  f.uid_off = Intern(event.involved_uid), f.uid_len = (uint32_t)event.involved_uid.size();
  events_.push_back(f);
}

Error Supervisor::Flush(std::vector<AppliedVerdict> *applied) {
  if (applied) applied->clear();
  stats_ = FlushStats{};
  last_events_.clear();
  last_runs_.clear();
  last_verdicts_.clear();
  last_pos_.clear();
  if (events_.empty()) return Error{};
  if (!ctx_) return Error{"supervisor not initialised"};
  const size_t n = events_.size(), n_runs = runs_.size();
  stats_.events = n, stats_.runs = n_runs;

  // ReadCheckpoint once per run (:264) — this and the writes below are the supervisor's real cost.
  double t0 = now_s();
  std::vector<std::optional<CheckpointedRequest>> row_of(n_runs);
  std::vector<uint8_t> run_stage(n_runs, 0), run_exists(n_runs, 0);
  for (size_t r = 0; r < n_runs; ++r) {
    const Run &run = runs_[r];
    if (!run.found) continue;
    row_of[r] = store_->ReadCheckpoint(HeapString(run.alg_off, run.alg_len), HeapString(run.req_off, run.req_len));
    if (row_of[r]) {
      run_exists[r] = 1;
      const auto it = stage_id_.find(row_of[r]->lifecycle_stage);
      run_stage[r] = it == stage_id_.end() ? 0 : it->second;
    }
  }
  double t1 = now_s();
  stats_.read_checkpoints_s = t1 - t0;

  // The records of a run made adjacent, runs in order of first arrival, records in arrival order: a
  // counting pass over the run indices (they were handed out in first-arrival order by OnEvent).
  std::vector<uint32_t> next(n_runs);
  uint32_t off = 0;
  for (size_t r = 0; r < n_runs; ++r) {
    next[r] = off;
    off += runs_[r].count;
  }
  col_kind_.resize(n), col_flags_.resize(n), col_stage_.resize(n), col_reason_.resize(n), col_key_.resize(n);
  src_.resize(n);
  last_pos_.resize(n);
  rows_.resize(n_runs);
  for (size_t i = 0; i < n; ++i) {
    const Flat &f = events_[i];
    const uint32_t j = next[f.run]++;
    col_kind_[j] = f.kind;
    col_reason_[j] = f.reason;
    col_flags_[j] = (uint8_t)(f.flags | (run_exists[f.run] ? NXS_FLAG_CHECKPOINT : 0));
    col_stage_[j] = run_stage[f.run];
    col_key_[j] = runs_[f.run].job_key;
    src_[j] = (uint32_t)i;
    last_pos_[i] = j;
  }
  double t2 = now_s();
  stats_.group_s = t2 - t1;

  uint64_t n_jobs = 0;
  if (keep_verdicts_) col_verdict_.resize(n);
  const nxs_batch in{col_kind_.data(), col_reason_.data(), col_flags_.data(), col_stage_.data(), col_key_.data(), n};
  const nxs_out out{keep_verdicts_ ? col_verdict_.data() : nullptr, rows_.data(), &n_jobs};
  const int rc = nxs_classify(ctx_, &in, &out);
  if (rc != NXS_OK) return Error{std::string("nxs_classify: ") + nxs_strerror(rc) + ": " + nxs_last_error(ctx_)};
  double t3 = now_s();
  stats_.classify_s = t3 - t2;
  stats_.job_rows = n_jobs;
  if (n_jobs != n_runs) return Error{"internal error: job rows and runs of the batch differ in number"};

  // The I/O half of superviseAction, once per run, from the job row (row r belongs to run r).
  for (uint64_t r = 0; r < n_jobs; ++r) {
    const nxs_job_row &row = rows_[r];
    const Run &run = runs_[r];
    if (!run.found || row.verdict == 0) continue; // unresolved events carry no run; nothing handled: nothing to do
    std::string trace;
    if (row.winner_idx != NXS_NO_WINNER) {
      const Flat &w = events_[src_[row.winner_idx]];
      trace = HeapString(w.msg_off, w.msg_len);
    }
    AppliedVerdict a;
    (void)Apply(HeapString(run.alg_off, run.alg_len), HeapString(run.req_off, run.req_len), row.verdict, trace, row_of[r], &a);
    if (applied && (a.deleted_job || a.upserted || a.skipped_finished || a.err)) applied->push_back(std::move(a));
  }
  stats_.apply_s = now_s() - t3;

  // keep the batch for AnalysisResultOf (arrival order)
  if (keep_verdicts_) {
    last_verdicts_.resize(n);
    for (size_t i = 0; i < n; ++i) last_verdicts_[i] = col_verdict_[last_pos_[i]];
  }
  last_events_.swap(events_);
  last_runs_.swap(runs_);
  last_heap_.swap(heap_);
  events_.clear();
  runs_.clear();
  heap_.clear();
  run_of_.clear();
  return Error{};
}

// The observable effects of superviseAction (:261-374) for one verdict.
Error Supervisor::Apply(const std::string &algorithm, const std::string &request_id, uint32_t verdict,
                        const std::string &trace, const std::optional<CheckpointedRequest> &checkpoint,
                        AppliedVerdict *out) {
  out->algorithm = algorithm;
  out->request_id = request_id;
  out->trace = trace;
  out->verdict = verdict;
  if (NXS_V_ERR(verdict) == NXS_ERR_NO_CHECKPOINT) { // :265-273: delete the job, report the read error
    (void)client_->DeleteJob(ns_, request_id);
    out->deleted_job = true;
    out->err = Error{"no checkpoint exists for the provided request, job will be deleted without metadata saved"};
    return out->err;
  }
  if (NXS_V_SKIP_FINISHED(verdict)) out->skipped_finished = true; // :276-279
  if (!NXS_V_UPSERT(verdict)) return Error{};
  if (NXS_V_DELETE_JOB(verdict)) { // :289, :314, :339
    if (!client_->DeleteJob(ns_, request_id)) {
      out->err = Error{"failed to delete an algorithm submission"}; // :292-295: no upsert, retry later
      return out->err;
    }
    out->deleted_job = true;
  }
  auto row = checkpoint; // checkpoint.DeepCopy() (:281)
  if (!row) {
    out->err = Error{"internal error: upsert verdict without a checkpoint"};
    return out->err;
  }
  row->lifecycle_stage = vocab_.stage_names[NXS_V_NEW_STAGE(verdict) & 7]; // :297, :324, :349, :362
  if (const char *fmt = nxs_cause_template(NXS_V_CAUSE_TPL(verdict))) {       // :298, :325, :350
    const char *msg = nxs_msg_template(NXS_V_MSG_TPL(verdict));
    std::string cause(strlen(fmt) + strlen(msg) + 1, '\0');
    cause.resize((size_t)snprintf(cause.data(), cause.size(), fmt, msg));
    row->algorithm_failure_cause = cause;
    row->algorithm_failure_details = trace; // :299, :326, :351
  }
  if (!store_->UpsertCheckpoint(*row)) { // :301, :328, :353, :364
    out->err = Error{"failed to update algorithm submission status"};
    return out->err;
  }
  out->upserted = true;
  return Error{};
}

// superviseAction for one result, decided on the host (same tables, no GPU round trip).
Error Supervisor::SuperviseAction(const RunStatusAnalysisResult &r) {
  uint32_t action = 0;
  for (uint32_t a = 1; a <= 4; ++a)
    if (r.Action == nxs_action_name(a)) action = a;
  const auto row = store_->ReadCheckpoint(r.Algorithm, r.RequestId); // :264
  AppliedVerdict out;
  if (!row) return Apply(r.Algorithm, r.RequestId, NXS_V_PACK(action, 0, 0, 0, 1, 0, 0, NXS_ERR_NO_CHECKPOINT), "", row, &out);
  uint32_t stage = 0;
  for (uint32_t s = 0; s < 8; ++s)
    if (row->lifecycle_stage == vocab_.stage_names[s]) stage = s;
  if ((vocab_.finished_mask >> stage) & 1u) return Error{}; // :276-279
  if (action == 0) return Error{"unknown analysis result action: " + r.Action}; // :371-372
  static const uint32_t new_stage[5] = {0, NXS_STAGE_SCHEDULING_FAILED, NXS_STAGE_FAILED, NXS_STAGE_DEADLINE_EXCEEDED,
                                        NXS_STAGE_RUNNING};
  static const uint32_t cause_tpl[5] = {0, 1, 2, 3, 0};
  uint32_t msg_tpl = 0;
  for (uint32_t m = 1; m < 7; ++m)
    if (r.RunStatusMessage == nxs_msg_template(m)) msg_tpl = m;
  const uint32_t v = NXS_V_PACK(action, new_stage[action], cause_tpl[action], msg_tpl, action != NXS_ACTION_RUNNING, 1, 0, 0);
  // a message outside the reference's own set is formatted as given
  if (cause_tpl[action] && msg_tpl == 0) {
    if (!client_->DeleteJob(ns_, r.RequestId)) return Error{"failed to delete an algorithm submission"};
    CheckpointedRequest c = *row;
    c.lifecycle_stage = vocab_.stage_names[new_stage[action]];
    const char *fmt = nxs_cause_template(cause_tpl[action]);
    std::string cause(strlen(fmt) + r.RunStatusMessage.size() + 1, '\0');
    cause.resize((size_t)snprintf(cause.data(), cause.size(), fmt, r.RunStatusMessage.c_str()));
    c.algorithm_failure_cause = cause;
    c.algorithm_failure_details = r.RunStatusTrace;
    return store_->UpsertCheckpoint(c) ? Error{} : Error{"failed to update algorithm submission status"};
  }
  return Apply(r.Algorithm, r.RequestId, v, r.RunStatusTrace, row, &out);
}

std::string FailureCause(uint32_t cause_tpl, uint32_t msg_tpl) {
  const char *fmt = nxs_cause_template(cause_tpl);
  if (!fmt) return std::string();
  const char *msg = nxs_msg_template(msg_tpl);
  std::string cause(strlen(fmt) + strlen(msg) + 1, '\0');
  cause.resize((size_t)snprintf(cause.data(), cause.size(), fmt, msg));
  return cause;
}

std::vector<CheckpointUpdate> EncodeCheckpointRows(const nxs_job_row *rows, uint64_t n_rows, const Vocabulary &vocab) {
  // the 4 x 7 possible cause strings are formatted once, not per row
  std::string causes[4][7];
  for (uint32_t c = 1; c < 4; ++c)
    for (uint32_t m = 0; m < 7; ++m) causes[c][m] = FailureCause(c, m);
  std::vector<CheckpointUpdate> out;
  for (uint64_t r = 0; r < n_rows; ++r) {
    const uint32_t v = rows[r].verdict;
    if (!NXS_V_UPSERT(v)) continue;
    CheckpointUpdate u;
    u.job_row = (uint32_t)r;
    u.lifecycle_stage = vocab.stage_names[NXS_V_NEW_STAGE(v) & 7];
    u.delete_job = NXS_V_DELETE_JOB(v) != 0;
    if (NXS_V_CAUSE_TPL(v)) {
      u.sets_failure = true;
      u.algorithm_failure_cause = causes[NXS_V_CAUSE_TPL(v)][NXS_V_MSG_TPL(v) % 7];
      u.details_record = rows[r].winner_idx;
    }
    out.push_back(std::move(u));
  }
  return out;
}

RunStatusAnalysisResult Supervisor::AnalysisResultOf(size_t i) const {
  RunStatusAnalysisResult r;
  if (i >= last_events_.size() || i >= last_verdicts_.size()) return r;
  const uint32_t v = last_verdicts_[i];
  if (NXS_V_ACTION(v) == 0) return r;
  const Flat &f = last_events_[i];
  const Run &run = last_runs_[f.run];
  r.Action = nxs_action_name(NXS_V_ACTION(v));
  r.RunStatusMessage = nxs_msg_template(NXS_V_MSG_TPL(v));
  r.RunStatusTrace = last_heap_.substr(f.msg_off, f.msg_len);
  r.ObjectUID = last_heap_.substr(f.uid_off, f.uid_len);
  r.ObjectKind = f.kind == NXS_KIND_JOB ? "Job" : "Pod"; // (only Job and Pod events are ever handled, :159,:210)
  r.RequestId = last_heap_.substr(run.req_off, run.req_len);
  r.Algorithm = last_heap_.substr(run.alg_off, run.alg_len);
  return r;
}

} // namespace nxs_host
