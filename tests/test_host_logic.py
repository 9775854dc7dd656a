"""CPU tests of the product's host-side pieces: the C-ABI library loads and exports every symbol the
header declares (no compute without a GPU), the host vocabulary matches the reference strings, and the
kernel's tile algorithm — walked through on the CPU with the product's own fold algebra
(tests/foldsim) — matches the oracle, including jobs that straddle chunk and tile boundaries."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import nexus_supervisor_b200 as nx
from oracle import cpu_oracle
from tests import util

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_functions():
    text = open(os.path.join(ROOT, "include", "nxs_classify.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(nxs_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = nx._abi.load()
    declared = _header_functions()
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/nxs_classify.h but not exported"
        assert name in nx._abi.SIGNATURES, f"{name} has no ctypes prototype"
    assert lib.nxs_abi_version() == 1


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(nx.NxsError) as e:
        nx.Classifier()
    assert e.value.status == nx._abi.NXS_ERR_NO_DEVICE


def test_vocabulary_matches_reference_strings():
    names = cpu_oracle.names()
    for i, r in enumerate(names["reason"]):
        assert nx.reason_name(i) == r
        if i:
            assert nx.intern_reason(r) == i and nx.REASON[r] == i
    assert nx.intern_reason("NoSuchReason") == 0 and nx.intern_reason("") == 0
    assert [nx.intern_kind(k) for k in ("Job", "Pod", "Node", "")] == [1, 2, 0, 0]

    for i, s in enumerate(names["stage"]):
        assert nx.stage_name(i) == s and nx.STAGE[s] == i
    assert [nx.action_name(i) for i in range(5)] == ["", "ToFailStuckInPending", "ToFailFatalError",
                                                    "ToFailDeadlineExceeded", "ToRunning"]
    # message / cause templates against the string-faithful oracle
    for kind, reason in (("Job", "FailedCreate"), ("Job", "DeadlineExceeded"), ("Job", "PodFailurePolicy"),
                         ("Pod", "Started"), ("Pod", "Failed"), ("Pod", "BackOff")):
        o = cpu_oracle.event_strings(kind, reason, "BUFFERED")
        cols = util.make_cols([nx.KIND[kind]], [nx.REASON[reason]], [7], [1], [1])
        v = nx.unpack_verdict(cpu_oracle.classify(cols)[0][0])
        assert nx.msg_template(v["msg_tpl"]) == o["message"]
        tpl = nx.cause_template(v["cause_tpl"])
        assert (tpl % o["message"] if tpl else "") == o["cause"]
    assert nx.FINISHED_MASK_DEFAULT == cpu_oracle.FINISHED_MASK_DEFAULT


def test_extension_container_reason_premap():
    """EXTENSION, no reference parity: the north-star's container-status vocabulary mapped to the events that
    carry it to the supervisor; classified, those events give what the reference gives for them (the oracle's
    verdict for the OOMKilled case is the reference test's own scenario: Job PodFailurePolicy -> FAILED)."""
    want = {"OOMKilled": ("Job", "PodFailurePolicy"), "ImagePullBackOff": ("Pod", "BackOff"),
            "CrashLoopBackOff": ("Pod", "BackOff"), "ErrImagePull": ("Pod", "Failed")}
    for name, (kind, reason) in want.items():
        assert nx.container_reason_to_event(name) == (nx.intern_kind(kind), nx.intern_reason(reason)), name
    for name in ("Completed", "ContainerCreating", "", "oomkilled"):
        assert nx.container_reason_to_event(name) is None
    k, r = nx.container_reason_to_event("OOMKilled")
    cols = util.make_cols([k], [r], [7], [2], [1])  # a RUNNING run
    _, rows = cpu_oracle.classify(cols)
    v = nx.unpack_verdict(int(rows["verdict"][0]))
    assert v["new_stage"] == nx.STAGE["FAILED"] and v["delete_job"] and v["upsert"]



def test_shard_of_is_splitmix_mod():
    for key in (0, 1, 0xDEADBEEF, 2**64 - 1):
        for g in (1, 2, 4, 8):
            assert nx.shard_of(key, g) == (cpu_oracle.splitmix64(key) % g if g > 1 else 0)


# ------------------------------------------------------------------ tile walk-through
_sim = None


def foldsim(cols, fm, tile, seed=1):
    global _sim
    if _sim is None:
        _sim = C.CDLL(os.path.join(ROOT, "tests", "foldsim", "_build", "libnxs_foldsim.so"))
    n = len(cols["kind"])
    v = np.zeros(n, np.uint32)
    rows = np.zeros(n, cpu_oracle.JOB_ROW)
    nj = C.c_uint64(0)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    rc = _sim.nxs_foldsim(p(cols["kind"]), p(cols["reason_id"]), p(cols["flags"]), p(cols["cur_stage"]),
                          p(cols["job_key"]), C.c_uint64(n), C.c_uint16(fm), C.c_uint32(tile), C.c_uint64(seed),
                          p(v), p(rows), C.byref(nj))
    assert rc == 0
    return v, rows[:nj.value]


@pytest.mark.parametrize("fm", util.MASKS)
def test_tile_algorithm_exhaustive_domain(fm):
    for per_job in (1, 2, 5):
        cols = util.exhaustive_cols(per_job)
        want = cpu_oracle.classify(cols, fm)
        got = foldsim(cols, fm, 1024)
        util.assert_same(*got, *want, f"per_job={per_job}")


@pytest.mark.parametrize("tile", [32, 64, 1024, 2048])
def test_tile_algorithm_random_jobs_cross_boundaries(tile):
    rng = np.random.default_rng(tile)
    for trial in range(40):
        n = int(rng.integers(1, 7000))
        max_len = int(rng.choice([1, 2, 3, 8, 40, 200, 3000]))
        cols = util.random_cols(rng, n, max_len, wide=bool(trial % 3))
        fm = int(rng.choice(util.MASKS + [int(rng.integers(0, 65536))]))
        want = cpu_oracle.classify(cols, fm)
        got = foldsim(cols, fm, tile, seed=trial)
        util.assert_same(*got, *want, f"trial={trial} n={n} max_len={max_len} fm={fm:#x}")


def test_tile_algorithm_synthetic_stream():
    for workload in (1, 2):
        cols = cpu_oracle.generate(workload, 100_000)
        want = cpu_oracle.classify(cols)
        got = foldsim(cols, util.FM_DEFAULT, 2048)
        util.assert_same(*got, *want, f"workload={workload}")


# ------------------------------------------------------------------ worker-range walk-through (lane per job)
def foldsim_ranges(cols, fm, range_records):
    global _sim
    if _sim is None:
        _sim = C.CDLL(os.path.join(ROOT, "tests", "foldsim", "_build", "libnxs_foldsim.so"))
    n = len(cols["kind"])
    v = np.zeros(n, np.uint32)
    rows = np.zeros(n, cpu_oracle.JOB_ROW)
    nj = C.c_uint64(0)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    rc = _sim.nxs_foldsim_ranges(p(cols["kind"]), p(cols["reason_id"]), p(cols["flags"]), p(cols["cur_stage"]),
                                 p(cols["job_key"]), C.c_uint64(n), C.c_uint16(fm), C.c_uint32(range_records),
                                 p(v), p(rows), C.byref(nj))
    assert rc == 0
    return v, rows[:nj.value]


@pytest.mark.parametrize("fm", util.MASKS)
def test_job_pass_exhaustive_domain(fm):
    """The job pass of the kernel (one lane per job: runs cut out of the ballot words by funnel shifts,
    summ_run_flags / summ_run_finish) over every (kind, reason, flags, stage) combination."""
    for per_job in (1, 2, 5, 33):
        cols = util.exhaustive_cols(per_job)
        util.assert_same(*foldsim_ranges(cols, fm, 128), *cpu_oracle.classify(cols, fm), f"per_job={per_job}")


@pytest.mark.parametrize("range_records", [32, 128, 256, 512])
def test_job_pass_random_jobs_cross_ranges(range_records):
    """Runs of 1..3000 records: inside a 32-record word, straddling words (funnel shift), longer than 32
    records (word by word) and straddling worker ranges (unres / suffix stitched with summ_combine)."""
    rng = np.random.default_rng(range_records)
    for trial in range(60):
        n = int(rng.integers(1, 9000))
        max_len = int(rng.choice([1, 2, 3, 8, 31, 32, 33, 40, 200, 3000]))
        cols = util.random_cols(rng, n, max_len, wide=bool(trial % 3))
        fm = int(rng.choice(util.MASKS + [int(rng.integers(0, 65536))]))
        want = cpu_oracle.classify(cols, fm)
        util.assert_same(*foldsim_ranges(cols, fm, range_records), *want,
                         f"trial={trial} n={n} max_len={max_len} fm={fm:#x}")


def test_job_pass_synthetic_stream():
    for workload in (1, 2):
        cols = cpu_oracle.generate(workload, 200_000)
        util.assert_same(*foldsim_ranges(cols, util.FM_DEFAULT, 512), *cpu_oracle.classify(cols), f"workload={workload}")


@pytest.mark.parametrize("shares", [0, 1, 2, 7, 32, 256, 257, 300])
def test_run_parallel_runs_every_share_exactly_once(shares):
    """The helper behind every host loop of the library (nxs_host.h): each share runs once, also when there
    are more shares than thread slots — the path a thread that cannot be started takes as well."""
    sim = C.CDLL(os.path.join(ROOT, "tests", "foldsim", "_build", "libnxs_foldsim.so"))
    sim.nxs_foldsim_run_parallel.restype = C.c_int
    sim.nxs_foldsim_run_parallel.argtypes = [C.c_uint, C.POINTER(C.c_ulonglong)]
    total = C.c_ulonglong(0)
    assert sim.nxs_foldsim_run_parallel(shares, C.byref(total)) == shares
    assert total.value == sum(t * t for t in range(shares))


@pytest.mark.parametrize("fm", util.MASKS)
def test_verdict_bytes_expand_to_the_oracle_verdicts(fm):
    """Batches of more than one slice bring the per-record verdicts back as one byte each; the host expands
    them with verdict_of_code().  Over the whole input domain: the byte built from the product's tables, then
    expanded, is the oracle's per-record verdict — and the byte is a multiple of four below 256."""
    sim = C.CDLL(os.path.join(ROOT, "tests", "foldsim", "_build", "libnxs_foldsim.so"))
    cols = util.exhaustive_cols(1)
    n = len(cols["kind"])
    code = np.zeros(n, np.uint8)
    got = np.zeros(n, np.uint32)
    sim.nxs_foldsim_verdict_via_code.restype = None
    sim.nxs_foldsim_verdict_via_code(C.c_void_p(cols["kind"].ctypes.data), C.c_void_p(cols["reason_id"].ctypes.data),
                                     C.c_void_p(cols["flags"].ctypes.data), C.c_void_p(cols["cur_stage"].ctypes.data),
                                     C.c_uint64(n), C.c_uint32(fm), C.c_void_p(code.ctypes.data), C.c_void_p(got.ctypes.data))
    want, _ = cpu_oracle.classify(cols, fm)
    bad = np.nonzero(got != want)[0]
    assert bad.size == 0, (bad[:5], got[bad[:5]], want[bad[:5]])
    assert not (code & 3).any()

