// supervisor_test.cpp — the reference's TestSupervisor (services/supervisor_test.go:542-580) against the
// C++ host mirror + the CUDA library.  Like the reference test it preloads the events with their Jobs /
// Pods, seeds the checkpoint store, lets the supervisor process everything and then reads each checkpoint
// back and compares its LifecycleStage — here with an in-memory store and a recording Kubernetes client
// instead of Scylla and the fake clientset (neither exists on this box).  It also asserts what the
// reference test never does: failure cause / details, job deletion, and the error paths.
//
// Usage: supervisor_test <scenarios.tsv>   (written from tests/golden/supervisor_test_scenarios.json)
// TSV columns: kind reason involved_name request_id algorithm seeded_stage expected_stage objects message
#include <cstdio>
#include <fstream>
#include <set>
#include <sstream>

#include "nxs_supervisor.h"

using namespace nxs_host;

namespace {
const char *kNexusComponentLabel = "app.kubernetes.io/component"; // models.NexusComponentLabel (unpinned value)
const char *kAlgorithmRun = "algorithm-run";                      // models.JobLabelAlgorithmRun (unpinned value)

struct FakeCache : ObjectCache { // the informer caches of the fake clientset
  std::map<std::pair<std::string, std::string>, CachedObject> objects;
  bool IsNexusRunEvent(const Event &e) override {
    const auto it = objects.find({e.involved_kind, e.involved_name});
    if (it == objects.end()) return true; // not decidable without the object: let the stale-event guard drop it
    const auto l = it->second.labels.find(kNexusComponentLabel);
    return l != it->second.labels.end() && l->second == kAlgorithmRun;
  }
  const CachedObject *GetCachedObject(const Event &e) override {
    const auto it = objects.find({e.involved_kind, e.involved_name});
    return it == objects.end() ? nullptr : &it->second;
  }
};
struct FakeStore : CqlStore { // nexus.checkpoints, PRIMARY KEY ((algorithm, id))
  std::map<std::pair<std::string, std::string>, CheckpointedRequest> rows;
  int reads = 0, upserts = 0;
  std::optional<CheckpointedRequest> ReadCheckpoint(const std::string &a, const std::string &id) override {
    ++reads;
    const auto it = rows.find({a, id});
    if (it == rows.end()) return std::nullopt;
    return it->second;
  }
  bool UpsertCheckpoint(const CheckpointedRequest &c) override {
    ++upserts;
    rows[{c.algorithm, c.id}] = c;
    return true;
  }
};
struct FakeKube : KubeClient {
  std::set<std::string> deleted;
  bool DeleteJob(const std::string &, const std::string &name) override {
    deleted.insert(name);
    return true;
  }
};
struct Scenario {
  std::string kind, reason, involved_name, request_id, algorithm, seeded, expected, objects, message;
};

int failures = 0;
#define CHECK(cond, ...)                 \
  do {                                   \
    if (!(cond)) {                       \
      ++failures;                        \
      printf("FAIL %s:%d: ", __FILE__, __LINE__); \
      printf(__VA_ARGS__);               \
      printf("\n");                      \
    }                                    \
  } while (0)
} // namespace

int main(int argc, char **argv) {
  if (argc < 2) return 2;
  std::vector<Scenario> scenarios;
  std::ifstream in(argv[1]);
  for (std::string line; std::getline(in, line);) {
    std::stringstream ss(line);
    Scenario s;
    std::getline(ss, s.kind, '\t'); std::getline(ss, s.reason, '\t'); std::getline(ss, s.involved_name, '\t');
    std::getline(ss, s.request_id, '\t'); std::getline(ss, s.algorithm, '\t'); std::getline(ss, s.seeded, '\t');
    std::getline(ss, s.expected, '\t'); std::getline(ss, s.objects, '\t'); std::getline(ss, s.message, '\t');
    if (!s.kind.empty()) scenarios.push_back(s);
  }
  CHECK(scenarios.size() == 8, "expected the 8 scenarios of TestSupervisor, got %zu", scenarios.size());

  Vocabulary vocab;
  FakeCache cache;
  FakeStore store;
  FakeKube kube;
  // k8sObjects = append(k8sObjects, get...Objects(id)...)  (supervisor_test.go:543-550) + the CQL seed
  for (const Scenario &s : scenarios) {
    const std::map<std::string, std::string> labels = {{kNexusComponentLabel, kAlgorithmRun},
                                                       {vocab.JobTemplateNameKey, s.algorithm}};
    if (s.objects.find("Job") != std::string::npos) cache.objects[{"Job", s.request_id}] = CachedObject{labels};
    if (s.objects.find("Pod") != std::string::npos) {
      auto pod = labels;
      pod[vocab.JobNameLabel] = s.request_id;
      cache.objects[{"Pod", s.involved_name}] = CachedObject{pod};
    }
    store.rows[{s.algorithm, s.request_id}] = CheckpointedRequest{s.algorithm, s.request_id, s.seeded, "", ""};
  }

  // f.supervisor.Init(f.ctx, &ProcessingConfig{100ms, 1s, 10, 10, 4})  (supervisor_test.go:553-559)
  Supervisor supervisor(&kube, "nexus", &store, &cache, 0, vocab);
  ProcessingConfig cfg;
  cfg.FailureRateBaseDelayMs = 100;
  cfg.FailureRateMaxDelayMs = 1000;
  cfg.RateLimitElementsPerSecond = 10;
  cfg.RateLimitElementsBurst = 10;
  cfg.Workers = 4;
  const Error init_err = supervisor.Init(cfg, 1 << 16);
  CHECK(!init_err, "supervisor init failed %s", init_err.message.c_str());
  if (init_err) return 1;

  // the informer delivers the preloaded events (AddFunc: c.onEvent)
  for (const Scenario &s : scenarios)
    supervisor.OnEvent(Event{s.kind, s.involved_name, "uid-" + s.involved_name, s.reason,
                             s.message.empty() ? "trace of " + s.reason : s.message});
  std::vector<AppliedVerdict> applied;
  const Error flush_err = supervisor.Flush(&applied);
  CHECK(!flush_err, "flush failed %s", flush_err.message.c_str());

  // validate...Objects(f, id, t): read the checkpoint back, compare LifecycleStage
  for (size_t i = 0; i < scenarios.size(); ++i) {
    const Scenario &s = scenarios[i];
    const auto result = store.ReadCheckpoint(s.algorithm, s.request_id);
    CHECK(result.has_value(), "cannot read a checkpoint %s", s.request_id.c_str());
    if (!result) continue;
    CHECK(result->lifecycle_stage == s.expected, "lifecycle stage for %s should be %s, but is %s", s.request_id.c_str(),
          s.expected.c_str(), result->lifecycle_stage.c_str());
    // beyond the reference test: cause / details / deletion
    const RunStatusAnalysisResult r = supervisor.AnalysisResultOf(i);
    const bool finished_before = s.seeded == "CANCELLED";
    if (finished_before) {
      CHECK(result->algorithm_failure_cause.empty() && !kube.deleted.count(s.request_id), "cancelled run was touched");
    } else if (r.Action == ToRunning) {
      CHECK(result->algorithm_failure_cause.empty() && !kube.deleted.count(s.request_id), "ToRunning must not delete");
    } else {
      CHECK(kube.deleted.count(s.request_id) == 1, "job %s should have been deleted", s.request_id.c_str());
      CHECK(result->algorithm_failure_details == "trace of " + s.reason, "details: %s", result->algorithm_failure_details.c_str());
      std::string want;
      if (r.Action == ToFailStuckInPending)
        want = "Algorithm submission was buffered, but failed to launch on the target cluster: " + r.RunStatusMessage;
      else if (r.Action == ToFailFatalError) want = "Algorithm encountered a fatal error during execution: " + r.RunStatusMessage;
      else want = r.RunStatusMessage;
      CHECK(result->algorithm_failure_cause == want, "cause: %s", result->algorithm_failure_cause.c_str());
    }
    CHECK(r.RequestId == s.request_id && r.Algorithm == s.algorithm && r.ObjectKind == s.kind, "analysis result ids");
  }
  CHECK(store.reads >= 8 && store.upserts == 7, "store traffic: %d reads, %d upserts", store.reads, store.upserts);

  auto row = [&](const char *a, const char *id) -> CheckpointedRequest & { return store.rows[std::make_pair(std::string(a), std::string(id))]; };

  // ---- paths the reference marks `coverage-ignore`
  // stale event: involved object missing from the cache (:161-164, :218-221) -> dropped
  supervisor.OnEvent(Event{"Job", "no-such-job", "u", "FailedCreate", "m"});
  // not a Nexus run (:155-157) -> dropped
  cache.objects[{"Job", "foreign"}] = CachedObject{{{"app", "other"}}};
  store.rows[{"", "foreign"}] = CheckpointedRequest{"", "foreign", "RUNNING", "", ""};
  supervisor.OnEvent(Event{"Job", "foreign", "u", "FailedCreate", "m"});
  // no checkpoint (:265-273) -> job deleted, error reported, nothing written
  cache.objects[{"Job", "orphan"}] = CachedObject{{{kNexusComponentLabel, kAlgorithmRun}, {vocab.JobTemplateNameKey, "alg"}}};
  supervisor.OnEvent(Event{"Job", "orphan", "u", "PodFailurePolicy", "m"});
  // two events of one run in one batch: the first failure wins, the later one is skipped
  store.rows[{"alg", "twice"}] = CheckpointedRequest{"alg", "twice", "RUNNING", "", ""};
  cache.objects[{"Job", "twice"}] = cache.objects[{"Job", "orphan"}];
  supervisor.OnEvent(Event{"Job", "twice", "u", "DeadlineExceeded", "first"});
  supervisor.OnEvent(Event{"Job", "twice", "u", "PodFailurePolicy", "second"});
  const int upserts_before = store.upserts;
  const Error e2 = supervisor.Flush(&applied);
  CHECK(!e2, "second flush failed %s", e2.message.c_str());
  CHECK(row("", "foreign").lifecycle_stage == "RUNNING" && !kube.deleted.count("foreign"), "foreign job touched");
  CHECK(!kube.deleted.count("no-such-job"), "stale event acted upon");
  CHECK(kube.deleted.count("orphan") == 1, "job without checkpoint must be deleted");
  bool orphan_err = false;
  for (const AppliedVerdict &a : applied)
    if (a.request_id == "orphan") orphan_err = (bool)a.err && !a.upserted;
  CHECK(orphan_err, "missing checkpoint must surface as an error");
  CHECK(row("alg", "twice").lifecycle_stage == "DEADLINE_EXCEEDED" && row("alg", "twice").algorithm_failure_details == "first",
        "first failure must win: %s / %s", row("alg", "twice").lifecycle_stage.c_str(),
        row("alg", "twice").algorithm_failure_details.c_str());
  CHECK(store.upserts == upserts_before + 1, "one upsert per run expected, got %d", store.upserts - upserts_before);

  // unknown action (:371-372)
  const Error unknown = supervisor.SuperviseAction(RunStatusAnalysisResult{"ToSomethingElse", "m", "t", "u", "Job", "twice2", "alg"});
  store.rows[{"alg", "twice2"}] = CheckpointedRequest{"alg", "twice2", "RUNNING", "", ""};
  const Error unknown2 = supervisor.SuperviseAction(RunStatusAnalysisResult{"ToSomethingElse", "m", "t", "u", "Job", "twice2", "alg"});
  CHECK((bool)unknown, "no checkpoint must be an error");
  CHECK(unknown2.message == "unknown analysis result action: ToSomethingElse", "got: %s", unknown2.message.c_str());
  const Error single = supervisor.SuperviseAction(RunStatusAnalysisResult{ToFailFatalError, "BackOff", "t", "u", "Pod", "twice2", "alg"});
  CHECK(!single && row("alg", "twice2").lifecycle_stage == "FAILED" &&
            row("alg", "twice2").algorithm_failure_cause == "Algorithm encountered a fatal error during execution: BackOff",
        "single-result path: %s", row("alg", "twice2").algorithm_failure_cause.c_str());

  // batched checkpoint-row encoder: byte-exact cause strings, one update per upserting row
  {
    const nxs_job_row rows[5] = {
        {1, NXS_V_PACK(NXS_ACTION_FATAL_ERROR, NXS_STAGE_FAILED, 2, 3, 1, 1, 0, 0), 7},          // Job/PodFailurePolicy
        {2, NXS_V_PACK(NXS_ACTION_RUNNING, NXS_STAGE_RUNNING, 0, 0, 0, 1, 0, 0), NXS_NO_WINNER}, // Pod/Started
        {3, NXS_V_PACK(0, 0, 0, 0, 0, 0, 1, 0), NXS_NO_WINNER},                                  // skipped: finished
        {4, NXS_V_PACK(0, 0, 0, 0, 1, 0, 0, 1), NXS_NO_WINNER},                                  // no checkpoint
        {5, NXS_V_PACK(NXS_ACTION_DEADLINE_EXCEEDED, NXS_STAGE_DEADLINE_EXCEEDED, 3, 2, 1, 1, 0, 0), 9}};
    const auto ups = EncodeCheckpointRows(rows, 5);
    CHECK(ups.size() == 3, "encoder: %zu updates", ups.size());
    if (ups.size() == 3) {
      CHECK(ups[0].lifecycle_stage == "FAILED" && ups[0].details_record == 7 && ups[0].delete_job &&
                ups[0].algorithm_failure_cause == "Algorithm encountered a fatal error during execution: Algorithm "
                                                  "encountered a fatal error during execution.",
            "encoder row 0: %s", ups[0].algorithm_failure_cause.c_str());
      CHECK(ups[1].lifecycle_stage == "RUNNING" && !ups[1].sets_failure && !ups[1].delete_job, "encoder row 1");
      CHECK(ups[2].job_row == 4 && ups[2].algorithm_failure_cause ==
                                       "Algorithm exceeded its max allowed run time limit or retry attempt count.",
            "encoder row 2: %s", ups[2].algorithm_failure_cause.c_str());
    }
  }

  printf(failures ? "FAILED (%d checks)\n" : "ok: TestSupervisor scenarios and error paths pass\n", failures);
  return failures ? 1 : 0;
}
