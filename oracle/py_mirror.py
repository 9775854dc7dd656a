"""Pure-Python mirror of the reference path, for cross-checking the C++ oracle on
small inputs.  Test infrastructure.  Written directly from
services/supervisor.go:137-374, independently of nxs_oracle.cpp: it works on the
string values the Go code switches on and only encodes the outcome at the end.
"""

KINDS = ["", "Job", "Pod"]
REASONS = ["", "FailedCreate", "DeadlineExceeded", "BackoffLimitExceeded", "PodFailurePolicy", "Started",
           "Failed", "BackOff", "Scheduled", "Pulling", "Pulled", "Created", "Killing", "SuccessfulCreate",
           "Completed", "FailedScheduling"]
STAGES = ["NEW", "BUFFERED", "RUNNING", "COMPLETED", "FAILED", "SCHEDULING_FAILED", "DEADLINE_EXCEEDED",
          "CANCELLED"] + ["STAGE_%d" % i for i in range(8, 16)]
ACTIONS = ["", "ToFailStuckInPending", "ToFailFatalError", "ToFailDeadlineExceeded", "ToRunning"]
MSG = ["", "Unable to launch a container for the algorithm - please review configuration and try again.",
       "Algorithm exceeded its max allowed run time limit or retry attempt count.",
       "Algorithm encountered a fatal error during execution.", "Started", "Failed", "BackOff"]
T0 = "Algorithm submission was buffered, but failed to launch on the target cluster: %s"
T1 = "Algorithm encountered a fatal error during execution: %s"


def on_event(kind, reason, supervised, found):
    """services/supervisor.go:137-259 -> (Action, RunStatusMessage) or None."""
    if not supervised:                      # :155
        return None
    result = None
    if kind == "Job":                       # :159
        if not found:                       # :161
            return None
        if reason == "FailedCreate":        # :172
            result = ("ToFailStuckInPending", MSG[1])
        elif reason in ("DeadlineExceeded", "BackoffLimitExceeded"):   # :183
            result = ("ToFailDeadlineExceeded", MSG[2])
        elif reason == "PodFailurePolicy":  # :194
            result = ("ToFailFatalError", MSG[3])
        else:                               # :205
            return None
    if kind == "Pod":                       # :210
        if not found:                       # :218
            return None
        if reason == "Started":             # :224
            result = ("ToRunning", reason)
        elif reason == "Failed":            # :234
            result = ("ToFailStuckInPending", reason)
        elif reason == "BackOff":           # :244
            result = ("ToFailFatalError", reason)
    return result


def supervise_action(action, message, row, finished):
    """services/supervisor.go:261-374.  row: dict(stage, cause) or None (no checkpoint).
    Returns dict(delete, upsert, skip, err); mutates row like the upsert would."""
    out = dict(delete=False, upsert=False, skip=False, err=0)
    if row is None:                         # :265-273
        out.update(delete=True, err=1)
        return out
    if finished(row["stage"]):              # :276
        out["skip"] = True
        return out
    if action == "ToFailStuckInPending":    # :284
        out.update(delete=True, upsert=True)
        row["stage"], row["cause"] = "SCHEDULING_FAILED", T0 % message
    elif action == "ToFailFatalError":      # :310
        out.update(delete=True, upsert=True)
        row["stage"], row["cause"] = "FAILED", T1 % message
    elif action == "ToFailDeadlineExceeded":  # :336
        out.update(delete=True, upsert=True)
        row["stage"], row["cause"] = "DEADLINE_EXCEEDED", message
    elif action == "ToRunning":             # :361
        out["upsert"] = True
        row["stage"] = "RUNNING"
    else:                                   # :371
        out["err"] = 2
    return out


def _pack(action, stage, cause, msg, delete, upsert, skip, err):
    return (action | stage << 3 | cause << 7 | msg << 9 | int(delete) << 12 | int(upsert) << 13 |
            int(skip) << 14 | err << 15)


def _cause_tpl(cause, message):
    return {T0 % message: 1, T1 % message: 2}.get(cause, 3 if cause == message else 0)


def classify(cols, finished_mask):
    """Returns (verdicts list, rows list of (job_key, verdict, winner_idx))."""
    def finished(stage):
        return bool((finished_mask >> STAGES.index(stage)) & 1)

    n = len(cols["kind"])
    verdicts, rows = [], []
    job = None
    for i in range(n):
        key = int(cols["job_key"][i])
        if i == 0 or key != int(cols["job_key"][i - 1]):
            job = dict(row=dict(stage=STAGES[int(cols["cur_stage"][i]) & 15], cause=None), last=0, cause_tpl=0,
                       msg=0, delete=False, upsert=False, skip=False, err=0, winner=0xFFFFFFFF)
        kind = int(cols["kind"][i])
        reason = int(cols["reason_id"][i])
        flags = int(cols["flags"][i])
        res = on_event(KINDS[kind] if kind < 3 else "Other", REASONS[reason] if reason < 16 else "Unknown",
                       bool(flags & 1), bool(flags & 2))
        v = 0
        if res is not None:
            action, message = res
            a, m = ACTIONS.index(action), MSG.index(message)
            exists = bool(flags & 4)
            own = dict(stage=STAGES[int(cols["cur_stage"][i]) & 15], cause=None) if exists else None
            o = supervise_action(action, message, own, finished)
            v = _pack(a, STAGES.index(own["stage"]) if o["upsert"] else 0,
                      _cause_tpl(own["cause"], message) if o["upsert"] and a != 4 else 0, m, o["delete"],
                      o["upsert"], o["skip"], o["err"])
            f = supervise_action(action, message, job["row"] if exists else None, finished)
            job["delete"] |= f["delete"]
            job["upsert"] |= f["upsert"]
            job["skip"] |= f["skip"]
            job["err"] = f["err"] or job["err"]
            if f["upsert"]:
                job["last"] = a
                if a != 4:
                    job["cause_tpl"], job["msg"], job["winner"] = _cause_tpl(job["row"]["cause"], message), m, i
        verdicts.append(v)
        if i + 1 == n or int(cols["job_key"][i + 1]) != key:
            stage_id = STAGES.index(job["row"]["stage"]) if job["upsert"] else 0
            rows.append((key, _pack(job["last"], stage_id, job["cause_tpl"], job["msg"], job["delete"],
                                    job["upsert"], job["skip"], job["err"]), job["winner"]))
    return verdicts, rows
