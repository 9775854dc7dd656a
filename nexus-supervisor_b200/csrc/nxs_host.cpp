// nxs_host.cpp — host-side vocabulary of the classification path: the interning
// dictionaries, the (kind, reason) decision table and the byte-exact message /
// cause templates.  Every entry cites the reference line it restates
// (services/supervisor.go in SneaksAndData/nexus-supervisor).
#include "nxs_host.h"

#include <cstring>

namespace {

struct SwitchArm {
  uint8_t kind;
  uint16_t reason;
  uint8_t action;
  uint8_t msg_tpl;
};
// The seven handled switch arms of onEvent.
constexpr SwitchArm kSwitchArms[] = {
    {NXS_KIND_JOB, NXS_REASON_FAILED_CREATE, NXS_ACTION_STUCK_IN_PENDING, 1},          // :172-182
    {NXS_KIND_JOB, NXS_REASON_DEADLINE_EXCEEDED, NXS_ACTION_DEADLINE_EXCEEDED, 2},     // :183-193
    {NXS_KIND_JOB, NXS_REASON_BACKOFF_LIMIT_EXCEEDED, NXS_ACTION_DEADLINE_EXCEEDED, 2}, // :183-193
    {NXS_KIND_JOB, NXS_REASON_POD_FAILURE_POLICY, NXS_ACTION_FATAL_ERROR, 3},          // :194-204
    {NXS_KIND_POD, NXS_REASON_STARTED, NXS_ACTION_RUNNING, 4},                         // :224-233
    {NXS_KIND_POD, NXS_REASON_FAILED, NXS_ACTION_STUCK_IN_PENDING, 5},                 // :234-243
    {NXS_KIND_POD, NXS_REASON_BACKOFF, NXS_ACTION_FATAL_ERROR, 6},                     // :244-253
};

const char *const kReasonNames[16] = {"",
                                      "FailedCreate",
                                      "DeadlineExceeded",
                                      "BackoffLimitExceeded",
                                      "PodFailurePolicy",
                                      "Started",
                                      "Failed",
                                      "BackOff",
                                      "Scheduled",
                                      "Pulling",
                                      "Pulled",
                                      "Created",
                                      "Killing",
                                      "SuccessfulCreate",
                                      "Completed",
                                      "FailedScheduling"};

// DecisionAction constants, :52-55
const char *const kActionNames[5] = {"", "ToFailStuckInPending", "ToFailFatalError", "ToFailDeadlineExceeded",
                                     "ToRunning"};

// LifecycleStage ids.  BUFFERED / RUNNING / CANCELLED are spelled as in
// test-resources/checkpoints.cql:35,44,98; the others are nexus-core constants
// (unpinned) — host-side data only, the kernel never sees strings.
const char *const kStageNames[8] = {"NEW",    "BUFFERED",          "RUNNING",           "COMPLETED",
                                    "FAILED", "SCHEDULING_FAILED", "DEADLINE_EXCEEDED", "CANCELLED"};

// RunStatusMessage per msg_tpl.  1..3 are the literals at :176,:187,:198;
// 4..6 are event.Reason itself (:227,:237,:247).
const char *const kMsgTemplates[7] = {
    "",
    "Unable to launch a container for the algorithm - please review configuration and try again.",
    "Algorithm exceeded its max allowed run time limit or retry attempt count.",
    "Algorithm encountered a fatal error during execution.",
    "Started",
    "Failed",
    "BackOff"};

// AlgorithmFailureCause formats per cause_tpl: :298, :325, :350 (verbatim message).
const char *const kCauseTemplates[4] = {
    nullptr, "Algorithm submission was buffered, but failed to launch on the target cluster: %s",
    "Algorithm encountered a fatal error during execution: %s", "%s"};

bool eq(const char *s, size_t len, const char *lit) { return strlen(lit) == len && memcmp(s, lit, len) == 0; }

} // namespace

namespace nxs {
void nxs_build_lut(uint8_t lut[128], uint16_t code[8]) {
  memset(lut, 0, 128);
  memset(code, 0, 8 * sizeof(uint16_t));
  // applier switch per action: new stage (:297, :324, :349, :362) and cause template (:298, :325, :350)
  static const uint8_t stage_of_action[5] = {0, NXS_STAGE_SCHEDULING_FAILED, NXS_STAGE_FAILED,
                                             NXS_STAGE_DEADLINE_EXCEEDED, NXS_STAGE_RUNNING};
  static const uint8_t cause_of_action[5] = {0, 1, 2, 3, 0};
  uint8_t arm = 0;
  for (const SwitchArm &a : kSwitchArms) {
    ++arm;
    lut[a.kind * 32 + a.reason] = arm;
    code[arm] = (uint16_t)(a.action | stage_of_action[a.action] << 3 | cause_of_action[a.action] << 7 | a.msg_tpl << 9);
  }
}
} // namespace nxs

extern "C" {

uint8_t nxs_intern_kind(const char *s, size_t len) {
  if (!s) return NXS_KIND_OTHER;
  if (eq(s, len, "Job")) return NXS_KIND_JOB; // :159
  if (eq(s, len, "Pod")) return NXS_KIND_POD; // :210
  return NXS_KIND_OTHER;
}

uint16_t nxs_intern_reason(const char *s, size_t len) {
  if (!s || len == 0) return NXS_REASON_UNKNOWN;
  for (uint16_t i = 1; i < 16; ++i)
    if (eq(s, len, kReasonNames[i])) return i;
  return NXS_REASON_UNKNOWN;
}

const char *nxs_reason_name(uint32_t id) { return id < 16 ? kReasonNames[id] : ""; }

// EXTENSION (see the header): container-status vocabulary -> the event that carries it to the reference.
int nxs_ext_container_reason_to_event(const char *s, size_t len, uint8_t *kind, uint16_t *reason_id) {
  uint8_t k = NXS_KIND_OTHER;
  uint16_t r = NXS_REASON_UNKNOWN;
  if (s) {
    if (eq(s, len, "OOMKilled")) k = NXS_KIND_JOB, r = NXS_REASON_POD_FAILURE_POLICY; // :311-312
    else if (eq(s, len, "ImagePullBackOff") || eq(s, len, "CrashLoopBackOff")) k = NXS_KIND_POD, r = NXS_REASON_BACKOFF;
    else if (eq(s, len, "ErrImagePull")) k = NXS_KIND_POD, r = NXS_REASON_FAILED;
  }
  if (kind) *kind = k;
  if (reason_id) *reason_id = r;
  return r != NXS_REASON_UNKNOWN;
}
const char *nxs_action_name(uint32_t a) { return a < 5 ? kActionNames[a] : ""; }
const char *nxs_stage_name(uint32_t s) { return s < 8 ? kStageNames[s] : ""; }
const char *nxs_msg_template(uint32_t i) { return i < 7 ? kMsgTemplates[i] : ""; }
const char *nxs_cause_template(uint32_t i) { return i < 4 ? kCauseTemplates[i] : nullptr; }

uint32_t nxs_shard_of(uint64_t job_key, uint32_t n_shards) {
  return n_shards <= 1 ? 0u : (uint32_t)(nxs::splitmix64(job_key) % n_shards);
}

} // extern "C"
