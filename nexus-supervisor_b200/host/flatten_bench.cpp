// flatten_bench.cpp — the host flattening layer (Supervisor::OnEvent / Flush) at scale: events in,
// verdicts applied.  Test and measurement driver, not part of the library.
//
//   flatten_bench <columns.bin> <rows_out.bin> [interleave]
//
// columns.bin holds a synthetic record stream as columns (u64 n, then kind u8[n], reason u16[n], flags
// u8[n], stage u8[n], key u64[n]); every record is turned into a Kubernetes-Event-shaped struct of
// STRINGS (kind, reason, involved name, uid, message) with a cache / store / client whose answers are
// what the columns say, and pushed through OnEvent; Flush classifies the batch on the GPU and applies
// the job rows to null sinks.  With `interleave` the two halves of the stream arrive alternately, so
// runs are not adjacent on arrival (per-run order is kept, as an informer keeps it).
// rows_out.bin: per job row that belongs to a resolved run, {u64 stream key, u32 verdict, u32 stream
// index of the winner event (0xFFFFFFFF: none)} — for the comparison with the string-mode oracle.
// stdout: one JSON line with the time of every stage.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <unordered_map>

#include "nxs_supervisor.h"

using namespace nxs_host;

namespace {
double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

struct Columns {
  uint64_t n = 0;
  std::vector<uint8_t> kind, flags, stage;
  std::vector<uint16_t> reason;
  std::vector<uint64_t> key;
};

// uid = "u<stream index>": the fakes read the record's flags back through it
uint64_t index_of(const std::string &uid) { return strtoull(uid.c_str() + 1, nullptr, 10); }

struct StreamCache : ObjectCache {
  const Columns *c;
  Vocabulary vocab;
  std::vector<CachedObject> objects; // one per record, built before the clock starts (the informer cache exists already)
  bool IsNexusRunEvent(const Event &e) override { return c->flags[index_of(e.involved_uid)] & NXS_FLAG_SUPERVISED; }
  const CachedObject *GetCachedObject(const Event &e) override {
    const uint64_t i = index_of(e.involved_uid);
    return (c->flags[i] & NXS_FLAG_FOUND) ? &objects[i] : nullptr;
  }
};
struct StreamStore : CqlStore { // answers from the columns; upserts are counted, not kept
  std::unordered_map<std::string, std::pair<uint8_t, bool>> rows; // request id -> (stage, exists)
  Vocabulary vocab;
  uint64_t reads = 0, upserts = 0;
  std::optional<CheckpointedRequest> ReadCheckpoint(const std::string &algorithm, const std::string &id) override {
    ++reads;
    const auto it = rows.find(id);
    if (it == rows.end() || !it->second.second) return std::nullopt;
    return CheckpointedRequest{algorithm, id, vocab.stage_names[it->second.first & 7], "", ""};
  }
  bool UpsertCheckpoint(const CheckpointedRequest &) override {
    ++upserts;
    return true;
  }
};
struct NullKube : KubeClient {
  uint64_t deletes = 0;
  bool DeleteJob(const std::string &, const std::string &) override {
    ++deletes;
    return true;
  }
};

std::string hex16(uint64_t v) {
  char b[20];
  snprintf(b, sizeof b, "%016llx", (unsigned long long)v);
  return b;
}
} // namespace

int main(int argc, char **argv) {
  if (argc < 3) return 2;
  const bool interleave = argc > 3 && !strcmp(argv[3], "interleave");
  Columns c;
  {
    std::ifstream in(argv[1], std::ios::binary);
    in.read(reinterpret_cast<char *>(&c.n), 8);
    c.kind.resize(c.n), c.reason.resize(c.n), c.flags.resize(c.n), c.stage.resize(c.n), c.key.resize(c.n);
    in.read(reinterpret_cast<char *>(c.kind.data()), c.n);
    in.read(reinterpret_cast<char *>(c.reason.data()), c.n * 2);
    in.read(reinterpret_cast<char *>(c.flags.data()), c.n);
    in.read(reinterpret_cast<char *>(c.stage.data()), c.n);
    in.read(reinterpret_cast<char *>(c.key.data()), c.n * 8);
    if (!in) return 3;
  }
  const uint64_t n = c.n;
  Vocabulary vocab;
  StreamCache cache;
  cache.c = &c;
  StreamStore store;
  NullKube kube;

  // ---- what exists before the supervisor sees anything: the events, the informer cache, the store
  static const char *kKind[4] = {"Node", "Job", "Pod", "Other"};
  std::vector<Event> events(n);
  cache.objects.resize(n);
  std::unordered_map<std::string, uint64_t> key_of_request;
  for (uint64_t i = 0; i < n; ++i) {
    const std::string request = hex16(c.key[i]);
    const uint8_t kind = c.kind[i] < 3 ? c.kind[i] : 3;
    const char *reason = nxs_reason_name(c.reason[i]);
    Event &e = events[i];
    e.involved_kind = kKind[kind];
    e.involved_name = kind == NXS_KIND_POD ? request + "-pod" : request;
    e.involved_uid = "u" + std::to_string(i);
    e.reason = (c.reason[i] != 0 && c.reason[i] < 16) ? reason : "Unknown" + std::to_string(c.reason[i]);
    e.message = "m" + std::to_string(i);
    cache.objects[i].labels = {{vocab.JobTemplateNameKey, "alg" + std::to_string(c.key[i] % 7)}, {vocab.JobNameLabel, request}};
    auto &row = store.rows[request];
    if (row.second == false) row = {c.stage[i], (c.flags[i] & NXS_FLAG_CHECKPOINT) != 0};
    key_of_request.emplace(request, c.key[i]);
  }
  // arrival order
  std::vector<uint64_t> order(n);
  if (interleave) {
    uint64_t half = n / 2;
    while (half < n && half > 0 && c.key[half] == c.key[half - 1]) ++half;
    uint64_t a = 0, b = half, k = 0;
    while (a < half || b < n) {
      if (a < half) order[k++] = a++;
      if (b < n) order[k++] = b++;
    }
  } else {
    for (uint64_t i = 0; i < n; ++i) order[i] = i;
  }

  Supervisor sup(&kube, "nexus", &store, &cache, 0, vocab);
  const Error init = sup.Init(ProcessingConfig(), n);
  if (init) {
    printf("%s\n", init.message.c_str());
    return 1;
  }
  sup.KeepRecordVerdicts(false);
  std::vector<AppliedVerdict> applied;
  double best_flatten = 1e30, best_total = 1e30;
  FlushStats st;
  for (int rep = 0; rep < 3; ++rep) { // (the first round also allocates the library's device staging)
    const double t0 = now_s();
    for (uint64_t k = 0; k < n; ++k) sup.OnEvent(events[order[k]]);
    const double t1 = now_s();
    const Error err = sup.Flush(&applied);
    const double t2 = now_s();
    if (err) {
      printf("%s\n", err.message.c_str());
      return 1;
    }
    if (t2 - t0 < best_total) best_total = t2 - t0, best_flatten = t1 - t0, st = sup.LastFlushStats();
  }

  // ---- what was applied, for the oracle comparison (one entry per resolved run with a verdict)
  {
    std::ofstream out(argv[2], std::ios::binary);
    for (const AppliedVerdict &a : applied) {
      const uint64_t key = key_of_request[a.request_id];
      const uint32_t v = a.verdict;
      uint32_t winner = NXS_NO_WINNER; // the winner's stream index travels in the trace ("m<index>")
      if (!a.trace.empty()) winner = (uint32_t)strtoull(a.trace.c_str() + 1, nullptr, 10);
      out.write(reinterpret_cast<const char *>(&key), 8);
      out.write(reinterpret_cast<const char *>(&v), 4);
      out.write(reinterpret_cast<const char *>(&winner), 4);
    }
  }
  printf("{\"events\": %llu, \"runs\": %llu, \"applied\": %zu, \"interleaved\": %s, \"flatten_s\": %.6f, "
         "\"read_checkpoints_s\": %.6f, \"group_s\": %.6f, \"classify_s\": %.6f, \"apply_s\": %.6f, \"total_s\": %.6f, "
         "\"events_per_s_flatten\": %.4g, \"events_per_s_total\": %.4g, \"store_reads\": %llu, \"upserts\": %llu, "
         "\"deletes\": %llu}\n",
         (unsigned long long)n, (unsigned long long)st.runs, applied.size(), interleave ? "true" : "false", best_flatten,
         st.read_checkpoints_s, st.group_s, st.classify_s, st.apply_s, best_total, n / best_flatten, n / best_total,
         (unsigned long long)store.reads, (unsigned long long)store.upserts, (unsigned long long)kube.deletes);
  return 0;
}
