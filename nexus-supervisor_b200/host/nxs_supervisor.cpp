// nxs_supervisor.cpp — see nxs_supervisor.h.  Citations: services/supervisor.go of the reference.
#include "nxs_supervisor.h"

#include <algorithm>
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
  const int rc = nxs_create(device_, max_batch, vocab_.finished_mask, &ctx_);
  if (rc != NXS_OK) return Error{std::string("nxs_create: ") + nxs_strerror(rc)};
  return Error{};
}

// onEvent, :137-259, minus the two switches (they run on the GPU).
void Supervisor::OnEvent(const Event &event) {
  Flat f;
  f.event = event;
  const bool supervised = cache_->IsNexusRunEvent(event); // :147
  if (supervised) f.flags |= NXS_FLAG_SUPERVISED;
  f.kind = nxs_intern_kind(event.involved_kind.data(), event.involved_kind.size());
  f.reason = nxs_intern_reason(event.reason.data(), event.reason.size());
  if (supervised && (f.kind == NXS_KIND_JOB || f.kind == NXS_KIND_POD)) {
    const auto obj = cache_->GetCachedObject(event.involved_kind, event.involved_name); // :160, :211
    if (obj) {
      f.flags |= NXS_FLAG_FOUND;
      auto label = [&](const std::string &key) {
        const auto it = obj->labels.find(key);
        return it == obj->labels.end() ? std::string() : it->second;
      };
      f.algorithm = label(vocab_.JobTemplateNameKey);                                              // :181, :232
      f.request_id = f.kind == NXS_KIND_JOB ? event.involved_name : label(vocab_.JobNameLabel);    // :180, :231
    }
  }
  // Events that cannot lead to a verdict still travel (the kernel drops them): they keep the batch
  // index == arrival order, and their job key groups them with nothing.
  f.job_key = (f.flags & NXS_FLAG_FOUND) ? JobKey(f.algorithm, f.request_id)
                                         : JobKey("\x01unresolved", event.involved_kind + "/" + event.involved_name);
  events_.push_back(std::move(f));
}

Error Supervisor::Flush(std::vector<AppliedVerdict> *applied) {
  if (applied) applied->clear();
  last_batch_.clear();
  last_verdicts_.clear();
  if (events_.empty()) return Error{};
  if (!ctx_) return Error{"supervisor not initialised"};
  const size_t n = events_.size();

  // Group the records of one run together, runs in order of first arrival, records in arrival order.
  std::unordered_map<uint64_t, size_t> first_seen;
  std::vector<size_t> order(n), rank(n);
  for (size_t i = 0; i < n; ++i) rank[i] = first_seen.emplace(events_[i].job_key, i).first->second;
  std::iota(order.begin(), order.end(), size_t(0));
  std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return rank[a] < rank[b]; });

  // ReadCheckpoint once per run (:264) — this and the writes below are the supervisor's real cost.
  std::vector<uint8_t> kind(n), flags(n), stage(n);
  std::vector<uint16_t> reason(n);
  std::vector<uint64_t> key(n);
  std::unordered_map<uint64_t, std::optional<CheckpointedRequest>> row_of; // job -> the row read at :264
  for (size_t j = 0; j < n; ++j) {
    const Flat &f = events_[order[j]];
    kind[j] = f.kind;
    reason[j] = f.reason;
    key[j] = f.job_key;
    uint8_t fl = f.flags, st = 0;
    if (f.flags & NXS_FLAG_FOUND) {
      auto it = row_of.find(f.job_key);
      if (it == row_of.end()) it = row_of.emplace(f.job_key, store_->ReadCheckpoint(f.algorithm, f.request_id)).first;
      if (it->second) {
        fl |= NXS_FLAG_CHECKPOINT;
        for (uint8_t s = 0; s < 8; ++s)
          if (it->second->lifecycle_stage == vocab_.stage_names[s]) st = s;
      }
    }
    flags[j] = fl;
    stage[j] = st;
  }

  std::vector<uint32_t> verdict(n);
  std::vector<nxs_job_row> rows(n);
  uint64_t n_jobs = 0;
  const nxs_batch in{kind.data(), reason.data(), flags.data(), stage.data(), key.data(), n};
  const nxs_out out{verdict.data(), rows.data(), &n_jobs};
  const int rc = nxs_classify(ctx_, &in, &out);
  if (rc != NXS_OK) return Error{std::string("nxs_classify: ") + nxs_strerror(rc) + ": " + nxs_last_error(ctx_)};

  // The I/O half of superviseAction, once per run, from the job row.
  size_t j = 0;
  for (uint64_t r = 0; r < n_jobs; ++r) {
    const nxs_job_row &row = rows[r];
    const size_t begin = j;
    while (j < n && key[j] == row.job_key) ++j; // the run's records are [begin, j)
    const Flat &first = events_[order[begin]];
    if (!(first.flags & NXS_FLAG_FOUND)) continue; // unresolved events carry no run
    const std::string trace = row.winner_idx != NXS_NO_WINNER ? events_[order[row.winner_idx]].event.message : "";
    AppliedVerdict a;
    (void)Apply(first.algorithm, first.request_id, row.verdict, trace, row_of[row.job_key], &a);
    if (applied && (a.deleted_job || a.upserted || a.skipped_finished || a.err)) applied->push_back(std::move(a));
  }
  // keep the batch for AnalysisResultOf (arrival order)
  last_verdicts_.assign(n, 0);
  for (size_t jj = 0; jj < n; ++jj) last_verdicts_[order[jj]] = verdict[jj];
  last_batch_.swap(events_);
  events_.clear();
  return Error{};
}

// The observable effects of superviseAction (:261-374) for one verdict.
Error Supervisor::Apply(const std::string &algorithm, const std::string &request_id, uint32_t verdict,
                        const std::string &trace, const std::optional<CheckpointedRequest> &checkpoint,
                        AppliedVerdict *out) {
  out->algorithm = algorithm;
  out->request_id = request_id;
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
  if (i >= last_batch_.size()) return r;
  const uint32_t v = last_verdicts_[i];
  if (NXS_V_ACTION(v) == 0) return r;
  const Flat &f = last_batch_[i];
  r.Action = nxs_action_name(NXS_V_ACTION(v));
  r.RunStatusMessage = nxs_msg_template(NXS_V_MSG_TPL(v));
  r.RunStatusTrace = f.event.message;
  r.ObjectUID = f.event.involved_uid;
  r.ObjectKind = f.event.involved_kind;
  r.RequestId = f.request_id;
  r.Algorithm = f.algorithm;
  return r;
}

} // namespace nxs_host
