"""Pins the oracle: the reference's own 8 known-answer scenarios
(services/supervisor_test.go:46-540 + test-resources/checkpoints.cql:31-101, extracted by
tests/golden/make_golden.py), the byte-exact strings of services/supervisor.go, and agreement between
the three independent CPU restatements (C++ string-faithful, C++ interned, pure Python)."""
import json
import os

import numpy as np
import pytest

from oracle import cpu_oracle, py_mirror
from tests import util

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "supervisor_test_scenarios.json")))

# models.LifecycleStage* constant -> stage spelling.  BUFFERED/RUNNING/CANCELLED are pinned by
# checkpoints.cql; the others are nexus-core values (unpinned) in the same convention.
STAGE_CONST = {"LifecycleStageSchedulingFailed": "SCHEDULING_FAILED", "LifecycleStageDeadlineExceeded":
               "DEADLINE_EXCEEDED", "LifecycleStageRunning": "RUNNING", "LifecycleStageFailed": "FAILED",
               "LifecycleStageCancelled": "CANCELLED"}


def test_golden_file_has_the_eight_scenarios():
    assert len(GOLDEN["scenarios"]) == 8
    assert GOLDEN["processing_config"] == {"RateLimitElementsPerSecond": 10, "RateLimitElementsBurst": 10,
                                           "Workers": 4}


@pytest.mark.parametrize("sc", GOLDEN["scenarios"], ids=lambda s: f"{s['kind']}-{s['reason']}-{s['seeded_stage']}")
def test_golden_scenario_string_mode(sc):
    out = cpu_oracle.event_strings(sc["kind"], sc["reason"], sc["seeded_stage"], message=sc["message"])
    assert out["stage"] == STAGE_CONST[sc["expected_stage_const"]]
    if sc["seeded_stage"] == "CANCELLED":
        assert out["skipped"] and not out["upsert"] and not out["delete_job"]
    else:
        assert out["upsert"]


@pytest.mark.parametrize("sc", GOLDEN["scenarios"], ids=lambda s: f"{s['kind']}-{s['reason']}-{s['seeded_stage']}")
def test_golden_scenario_interned_mode(sc):
    names = cpu_oracle.names()
    cols = util.make_cols([names["kind"].index(sc["kind"])], [names["reason"].index(sc["reason"])], [7],
                          [names["stage"].index(sc["seeded_stage"])], [12345])
    for mode in ("interned", "strings"):
        verdict, rows = cpu_oracle.classify(cols, mode=mode)
        v = int(verdict[0])
        want = names["stage"].index(STAGE_CONST[sc["expected_stage_const"]])
        final = (v >> 3) & 15 if (v >> 13) & 1 else int(cols["cur_stage"][0])
        assert final == want
        assert len(rows) == 1 and int(rows[0]["verdict"]) & ~(7 | 7 << 9) == v & ~(7 | 7 << 9)


def test_strings_are_byte_exact():
    """services/supervisor.go:176,187,198 (messages) and :298,:325,:350 (cause formats)."""
    o = cpu_oracle.event_strings("Job", "FailedCreate", "BUFFERED", message="trace-1")
    assert o["action"] == "ToFailStuckInPending"
    assert o["message"] == "Unable to launch a container for the algorithm - please review configuration and try again."
    assert o["cause"] == ("Algorithm submission was buffered, but failed to launch on the target cluster: "
                          + o["message"])
    assert o["details"] == "trace-1" and o["stage"] == "SCHEDULING_FAILED" and o["delete_job"]
    o = cpu_oracle.event_strings("Job", "BackoffLimitExceeded", "RUNNING")
    assert o["action"] == "ToFailDeadlineExceeded"
    assert o["message"] == "Algorithm exceeded its max allowed run time limit or retry attempt count."
    assert o["cause"] == o["message"]
    o = cpu_oracle.event_strings("Job", "PodFailurePolicy", "RUNNING")
    assert o["cause"] == ("Algorithm encountered a fatal error during execution: "
                          "Algorithm encountered a fatal error during execution.")
    o = cpu_oracle.event_strings("Pod", "BackOff", "BUFFERED")
    assert o["action"] == "ToFailFatalError" and o["message"] == "BackOff"
    assert o["cause"] == "Algorithm encountered a fatal error during execution: BackOff"
    o = cpu_oracle.event_strings("Pod", "Failed", "BUFFERED")
    assert o["cause"] == "Algorithm submission was buffered, but failed to launch on the target cluster: Failed"
    o = cpu_oracle.event_strings("Pod", "Started", "BUFFERED")
    assert o["action"] == "ToRunning" and o["stage"] == "RUNNING" and o["cause"] == "" and not o["delete_job"]


def test_guards():
    """:155 not supervised, :161/:218 stale event, :205/:254 default arms, :265 no checkpoint."""
    assert not cpu_oracle.event_strings("Job", "FailedCreate", "BUFFERED", supervised=False)["produced"]
    assert not cpu_oracle.event_strings("Job", "FailedCreate", "BUFFERED", found=False)["produced"]
    assert not cpu_oracle.event_strings("Pod", "BackOff", "BUFFERED", found=False)["produced"]
    assert not cpu_oracle.event_strings("Job", "Started", "BUFFERED")["produced"]
    assert not cpu_oracle.event_strings("Pod", "FailedCreate", "BUFFERED")["produced"]
    assert not cpu_oracle.event_strings("Node", "Failed", "BUFFERED")["produced"]
    o = cpu_oracle.event_strings("Pod", "BackOff", "BUFFERED", checkpoint=False)
    assert o["produced"] and o["err"] == 1 and o["delete_job"] and not o["upsert"] and o["stage"] == "BUFFERED"


def test_unknown_action_is_an_error():
    """services/supervisor.go:371-372."""
    assert cpu_oracle.apply_action_string("ToSomethingElse") == 2
    assert cpu_oracle.apply_action_string("ToRunning") == 0


@pytest.mark.parametrize("fm", util.MASKS)
def test_three_restatements_agree_exhaustively(fm):
    for per_job in (1, 3):
        cols = util.exhaustive_cols(per_job)
        v_i, r_i = cpu_oracle.classify(cols, fm, mode="interned")
        v_s, r_s = cpu_oracle.classify(cols, fm, mode="strings")
        util.assert_same(v_s, r_s, v_i, r_i, "strings vs interned")
    small = {k: v[::7][:4000] for k, v in util.exhaustive_cols(2).items()}
    v_i, r_i = cpu_oracle.classify(small, fm)
    v_p, r_p = py_mirror.classify(small, fm)
    assert np.array_equal(np.array(v_p, np.uint32), v_i)
    assert r_p == [tuple(int(x) for x in r) for r in r_i]


def test_random_jobs_three_way_and_threads():
    rng = np.random.default_rng(7)
    for max_len in (1, 4, 50):
        cols = util.random_cols(rng, 6000, max_len)
        for fm in (util.FM_DEFAULT, 0, 0x5A5A):
            v_i, r_i = cpu_oracle.classify(cols, fm)
            v_s, r_s = cpu_oracle.classify(cols, fm, mode="strings")
            v_t, r_t = cpu_oracle.classify(cols, fm, threads=3)
            util.assert_same(v_s, r_s, v_i, r_i, "strings vs interned")
            util.assert_same(v_t, r_t, v_i, r_i, "threads vs single")
        sub = {k: v[:1500] for k, v in cols.items()}
        v_i, r_i = cpu_oracle.classify(sub, util.FM_DEFAULT)
        v_p, r_p = py_mirror.classify(sub, util.FM_DEFAULT)
        assert np.array_equal(np.array(v_p, np.uint32), v_i)
        assert r_p == [tuple(int(x) for x in r) for r in r_i]


def test_default_mask_closed_form():
    """Under the default mask: the first failure record wins if the row was not finished, else
    RUNNING if any Started, else unchanged."""
    cols = cpu_oracle.generate(2, 30000)
    verdict, rows = cpu_oracle.classify(cols)
    keys = cols["job_key"]
    starts = np.nonzero(np.r_[True, keys[1:] != keys[:-1]])[0]
    ends = np.r_[starts[1:], len(keys)]
    assert len(rows) == len(starts)
    for (lo, hi), row in list(zip(zip(starts, ends), rows))[:4000]:
        acts = verdict[lo:hi] & 7
        he = (acts != 0) & ((cols["flags"][lo:hi] & 4) != 0)
        fin0 = (util.FM_DEFAULT >> int(cols["cur_stage"][lo])) & 1
        rv = int(row["verdict"])
        if fin0 or not he.any():
            assert (rv >> 13) & 1 == 0 and int(row["winner_idx"]) == 0xFFFFFFFF
            continue
        fails = np.nonzero(he & (acts != 4))[0]
        if fails.size:
            assert int(row["winner_idx"]) == lo + fails[0]
            assert (rv >> 3) & 15 == {1: 5, 2: 4, 3: 6}[int(acts[fails[0]])]
        else:
            assert (rv >> 3) & 15 == 2 and int(row["winner_idx"]) == 0xFFFFFFFF


def test_generator_statistics_and_shards():
    cols = cpu_oracle.generate(2, 200000)
    n = len(cols["kind"])
    keys = cols["job_key"]
    n_jobs = 1 + int((keys[1:] != keys[:-1]).sum())
    assert 2.2 < n / n_jobs < 2.6                      # 1 + Geometric(0.4), capped at 8
    handled_fail = np.isin(cols["reason_id"], [1, 2, 3, 4, 6, 7]).mean()
    assert 0.09 < handled_fail < 0.11                  # 10 % failures
    assert 0.28 < (cols["reason_id"] == 5).mean() < 0.32
    runs = np.diff(np.nonzero(np.r_[True, keys[1:] != keys[:-1], True])[0])
    assert runs.max() <= 8
    # shards partition the stream and keep jobs whole
    parts = [cpu_oracle.generate(2, 200000, shard=s, n_shards=4) for s in range(4)]
    assert sum(len(p["kind"]) for p in parts) == n
    for s, p in enumerate(parts):
        sm = np.array([cpu_oracle.splitmix64(int(k)) % 4 for k in p["job_key"][:200]])
        assert (sm == s).all()
    u = cpu_oracle.generate(1, 100000)
    assert (u["flags"] == 7).all() and set(np.unique(u["cur_stage"])) == {1, 2}
    assert abs((u["reason_id"] == 8).mean() - 0.125) < 0.01
