"""GPU parity tests: the CUDA path, called through the C ABI, against the CPU oracle — bit-exact.
Run on the B200 box with `pytest -m gpu`."""
import json
import os

import numpy as np
import pytest

import nexus_supervisor_b200 as nx
from oracle import cpu_oracle
from tests import util

pytestmark = pytest.mark.gpu

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "supervisor_test_scenarios.json")))
STAGE_CONST = {"LifecycleStageSchedulingFailed": "SCHEDULING_FAILED", "LifecycleStageDeadlineExceeded":
               "DEADLINE_EXCEEDED", "LifecycleStageRunning": "RUNNING", "LifecycleStageFailed": "FAILED",
               "LifecycleStageCancelled": "CANCELLED"}


@pytest.fixture(scope="module")
def clf():
    with nx.Classifier(device=0, max_records=1 << 23) as c:
        yield c


def test_golden_scenarios(clf):
    """The 8 scenarios of services/supervisor_test.go, one batch, one record per job."""
    sc = GOLDEN["scenarios"]
    cols = util.make_cols([nx.KIND[s["kind"]] for s in sc], [nx.REASON[s["reason"]] for s in sc], [7] * len(sc),
                          [nx.STAGE[s["seeded_stage"]] for s in sc], range(100, 100 + len(sc)))
    verdict, rows = clf.classify(cols)
    assert len(rows) == len(sc)
    for s, v, row in zip(sc, verdict, rows):
        f = nx.unpack_verdict(v)
        final = f["new_stage"] if f["upsert"] else nx.STAGE[s["seeded_stage"]]
        assert nx.stage_name(final) == STAGE_CONST[s["expected_stage_const"]], s
        assert nx.unpack_verdict(row["verdict"])["new_stage"] == f["new_stage"]
    util.assert_same(verdict, rows, *cpu_oracle.classify(cols), "golden")
    # the OOMKilled scenario of BASELINE config 0: FAILED row, T1 + M2 cause, job deleted
    g5 = nx.unpack_verdict(verdict[4])
    assert (g5["new_stage"], g5["cause_tpl"], g5["msg_tpl"], g5["delete_job"]) == (nx.STAGE["FAILED"], 2, 3, 1)


@pytest.mark.parametrize("fm", util.MASKS)
def test_exhaustive_domain(fm):
    with nx.Classifier(device=0, max_records=1 << 18, finished_mask=fm) as c:
        for per_job in (1, 3):
            cols = util.exhaustive_cols(per_job)
            util.assert_same(*c.classify(cols), *cpu_oracle.classify(cols, fm), f"fm={fm:#x} per_job={per_job}")


@pytest.mark.parametrize("n", [1, 2, 31, 32, 33, 2047, 2048, 2049, 4095, 4097, 6144, 8191, 8192, 8193, 16385, 24576,
                               100_003, 1_000_000])
def test_ragged_sizes(clf, n):
    rng = np.random.default_rng(n)
    cols = util.random_cols(rng, n, 5)
    util.assert_same(*clf.classify(cols), *cpu_oracle.classify(cols), f"n={n}")


def test_empty_batch(clf):
    verdict, rows = clf.classify(util.make_cols([], [], [], [], []))
    assert len(verdict) == 0 and len(rows) == 0


@pytest.mark.parametrize("max_len", [1, 2, 8, 40, 700, 5000, 400_000])
def test_jobs_straddling_warps_tiles_and_ctas(clf, max_len):
    """Jobs longer than a warp chunk, a tile and many tiles (look-back across head-less tiles)."""
    rng = np.random.default_rng(max_len)
    n = 700_001
    for fm in (util.FM_DEFAULT, 0):
        cols = util.random_cols(rng, n, max_len, wide=(fm == 0))
        with nx.Classifier(device=0, max_records=n, finished_mask=fm) as c:
            util.assert_same(*c.classify(cols), *cpu_oracle.classify(cols, fm), f"max_len={max_len} fm={fm:#x}")


@pytest.mark.parametrize("length", [31, 32, 33, 63, 64, 65, 511, 512, 513, 8191, 8192, 8193])
def test_jobs_of_exact_boundary_lengths(clf, length):
    """Jobs whose length is exactly a chunk (32), a worker range (512) or a tile (8192), one less and one
    more, at every alignment: the stream starts with a job of 0..length-1 records, and a few single-record
    jobs are sprinkled in so that the alignment drifts.  Dense and sparse records (mostly handled failures,
    mostly ignored events) so that stops fall on the first, the last and no record of a run."""
    rng = np.random.default_rng(length)
    n = 600_000 if length < 1000 else 1_500_000
    for dense in (True, False):
        for lead in (0, 1, length // 2, length - 1):
            lens = [lead] if lead else []
            while sum(lens) < n:
                lens.append(length)
                if rng.random() < 0.05:
                    lens.append(1)
            lens = np.array(lens)
            with np.errstate(over="ignore"):
                ids = np.arange(1, len(lens) + 1, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15)
            keys = np.repeat(ids, lens)[:n]
            if dense:
                kind, reason = rng.integers(1, 3, n), rng.integers(1, 8, n)
            else:
                kind = rng.integers(1, 3, n)
                reason = np.where(rng.random(n) < 0.002, rng.integers(1, 8, n), 9)
            stage = np.repeat(rng.integers(1, 3, len(lens)), lens)[:n]
            cols = util.make_cols(kind, reason, np.full(n, 7), stage, keys)
            for fm in (util.FM_DEFAULT, 0):
                if fm == util.FM_DEFAULT:
                    util.assert_same(*clf.classify(cols), *cpu_oracle.classify(cols), f"length={length} lead={lead} dense={dense}")
                else:
                    with nx.Classifier(device=0, max_records=n, finished_mask=0) as c:
                        util.assert_same(*c.classify(cols), *cpu_oracle.classify(cols, 0), f"length={length} lead={lead} dense={dense} fm=0")


def test_single_job_batch(clf):
    n = 300_000
    rng = np.random.default_rng(3)
    cols = util.random_cols(rng, n, 1, wide=False)
    cols["job_key"][:] = 42
    for fm in (util.FM_DEFAULT, 0):
        with nx.Classifier(device=0, max_records=n, finished_mask=fm) as c:
            verdict, rows = c.classify(cols)
            assert len(rows) == 1
            util.assert_same(verdict, rows, *cpu_oracle.classify(cols, fm), f"single job fm={fm:#x}")


@pytest.mark.parametrize("tile", ["2048", "4096", "8192"])
def test_tile_variants_and_context_reuse(tile, monkeypatch):
    monkeypatch.setenv("NXS_TILE", tile)
    rng = np.random.default_rng(int(tile))
    with nx.Classifier(device=0, max_records=2_000_000) as c:
        for n, max_len in ((1_500_000, 6), (5, 2), (777_777, 3000), (2_000_000, 2), (4096, 1), (8192, 1)):
            cols = util.random_cols(rng, n, max_len)
            util.assert_same(*c.classify(cols), *cpu_oracle.classify(cols), f"tile={tile} n={n}")
        assert c.launches == 6


def test_soak_many_launches_random_sizes(clf):
    """Many launches on one context with sizes around every hand-over of the pipeline (one tile, a few tiles,
    one tile per CTA, a little more than one per CTA and stage), long and short jobs: a rare ordering
    problem between the producer, the counter warps and the workers shows up as a wrong row or a trap."""
    rng = np.random.default_rng(20261107)
    sizes = [1, 8191, 8192, 8193, 16384, 16385]
    sizes += [int(x) for x in rng.integers(1, 60_000, 40)]
    sizes += [int(x) for x in rng.integers(148 * 8192 - 9000, 148 * 8192 + 9000, 10)]
    sizes += [int(x) for x in rng.integers(2 * 148 * 8192 - 9000, 2 * 148 * 8192 + 20_000, 10)]
    sizes += [int(x) for x in rng.integers(100_000, 5_000_000, 40)]
    for k, n in enumerate(sizes):
        max_len = (1, 2, 5, 40, 5000)[k % 5]
        cols = util.random_cols(rng, n, max_len)
        util.assert_same(*clf.classify(cols), *cpu_oracle.classify(cols), f"soak launch {k} n={n} max_len={max_len}")


def test_descriptor_generation_wraps():
    """Tile descriptors carry a 7-bit launch generation instead of being cleared before every launch; after
    127 launches of one context the scratch is zeroed and the count starts over.  300 launches of changing
    size on one context: descriptors of every earlier generation lie around while each launch runs."""
    rng = np.random.default_rng(127)
    with nx.Classifier(device=0, max_records=400_000) as c:
        for k in range(300):
            n = int(rng.integers(1, 400_000)) if k % 3 else int(rng.integers(1, 20_000))
            cols = util.random_cols(rng, n, (1, 3, 40, 3000)[k % 4])
            util.assert_same(*c.classify(cols), *cpu_oracle.classify(cols), f"launch {k} n={n}")
        assert c.launches == 300


def test_host_path_slices_on_job_boundaries():
    """The host-pointer entry point processes large batches in ~8M-record slices (upload, classify and
    download overlapped); slices are cut on job boundaries and winner indices stay batch-relative."""
    rng = np.random.default_rng(99)
    n = 20_000_003
    cols = util.random_cols(rng, n, 60, wide=False)
    with nx.Classifier(device=0, max_records=n) as c:
        for _ in range(2):  # the second call reuses the staging buffers and events
            util.assert_same(*c.classify(cols), *cpu_oracle.classify(cols, threads=cpu_oracle.hardware_threads()),
                             "sliced host path")
        assert c.launches == 2 * 3 * 2  # per slice: the expansion of the packed small columns, then the classification
        up, down = c.last_wire_bytes
        assert up == 10 * n and n + 16 * len(cols['kind']) // 60 <= down < 4 * n + 16 * n  # (verdicts travel as bytes)


@pytest.mark.parametrize("n", [(8 << 20) - 1, 8 << 20, (8 << 20) + 1, (9 << 20), (9 << 20) + 1, (16 << 20) + 5])
def test_host_path_sizes_around_the_slice_cuts(n):
    """Batches whose size sits on the edges of the slicing rule (one slice of up to 9 Mi records, otherwise
    8 Mi-record slices and a last one of up to 9 Mi), with a job lying across every nominal cut."""
    rng = np.random.default_rng(n & 0xFFFF)
    cols = util.random_cols(rng, n, 3, wide=False)
    for cut in range(8 << 20, n, 8 << 20):
        cols["job_key"][cut - 700:min(n, cut + 900)] = np.uint64(0xC0FFEE00 + cut)
    want = cpu_oracle.classify(cols, threads=cpu_oracle.hardware_threads())
    with nx.Classifier(device=0, max_records=n) as c:
        util.assert_same(*c.classify(cols), *want, f"n={n}")
        _, rows = c.classify(cols, want_verdict=False)
        assert np.array_equal(rows, want[1])


def test_host_path_error_after_enqueued_slices_leaves_nothing_in_flight():
    """A job of four million records straddling a slice cut makes nxs_classify give up with
    NXS_ERR_CAPACITY after earlier slices have been uploaded, classified and queued for download.  The call
    must drain its copy streams before it returns: the caller frees (here: poisons and drops) its buffers
    right away, and the same context then classifies another batch correctly."""
    import ctypes as C
    rng = np.random.default_rng(7)
    n = 20_000_000
    cols = util.random_cols(rng, n, 5, wide=False)
    cols["job_key"][15_500_000:19_500_000] = np.uint64(0xABCDEF0123456789)  # one job across the second cut (2 x 8Mi)
    lib = nx._abi.load()
    with nx.Classifier(device=0, max_records=n) as c:
        def pinned(nbytes):
            p = C.c_void_p()
            assert lib.nxs_alloc_host(nbytes, C.byref(p)) == 0
            return p
        bufs = {k: pinned(v.nbytes) for k, v in cols.items()}
        for k, v in cols.items():
            C.memmove(bufs[k], v.ctypes.data, v.nbytes)
        out_v, out_r = pinned(4 * n), pinned(16 * n)
        nj = C.c_uint64(0)
        batch = nx._abi.Batch(bufs["kind"], bufs["reason_id"], bufs["flags"], bufs["cur_stage"], bufs["job_key"], n)
        out = nx._abi.Out(out_v, out_r, C.addressof(nj))
        rc = lib.nxs_classify(c._ctx, C.byref(batch), C.byref(out))
        assert rc == nx._abi.NXS_ERR_CAPACITY, lib.nxs_strerror(rc)
        assert b"single job" in lib.nxs_last_error(c._ctx)
        assert c.launches >= 1  # at least one slice had been enqueued before the call gave up
        for p in list(bufs.values()) + [out_v, out_r]:  # the caller's buffers are its own again: free them now
            lib.nxs_free_host(p)
        small = util.random_cols(rng, 3_000_000, 40)
        util.assert_same(*c.classify(small), *cpu_oracle.classify(small, threads=4), "after the failed call")
        # the same failure with ordinary pageable buffers (bounce buffers and their helper threads in play)
        launched = c.launches
        with pytest.raises(nx.NxsError) as e:
            c.classify(cols)
        assert e.value.status == nx._abi.NXS_ERR_CAPACITY and c.launches > launched
        util.assert_same(*c.classify(small), *cpu_oracle.classify(small, threads=4), "after the failed pageable call")
        # the whole batch still goes through the device-resident entry point, as the error text says
        import torch
        dev = {k: torch.from_numpy(v.view(np.uint8).copy()).cuda() for k, v in cols.items()}
        v_d = torch.empty(n, dtype=torch.int32, device="cuda")
        r_d = torch.empty((n, 2), dtype=torch.int64, device="cuda")
        n_d = torch.zeros(1, dtype=torch.int64, device="cuda")
        c.classify_device(n, dev["kind"].data_ptr(), dev["reason_id"].data_ptr(), dev["flags"].data_ptr(),
                          dev["cur_stage"].data_ptr(), dev["job_key"].data_ptr(), v_d.data_ptr(), r_d.data_ptr(),
                          n_d.data_ptr())
        c.sync()
        want_v, want_r = cpu_oracle.classify(cols, threads=cpu_oracle.hardware_threads())
        jobs = int(n_d.item())
        got_r = r_d[:jobs].cpu().numpy().view(np.uint64).reshape(-1, 2).copy().view(cpu_oracle.JOB_ROW).reshape(-1)
        util.assert_same(v_d.cpu().numpy().view(np.uint32), got_r, want_v, want_r, "device path, 4M-record job")


def test_two_contexts_launching_concurrently():
    """Two contexts on one GPU, each on its own stream, launched back to back without synchronising in
    between: the kernels share the SMs, each context has its own tile counter and descriptors."""
    import torch
    n = 6_000_000
    rng = np.random.default_rng(17)
    runs = []
    for i in range(2):
        cols = util.random_cols(rng, n, 4 + 20 * i)
        stream = torch.cuda.Stream()
        clf = nx.Classifier(device=0, max_records=n, stream=stream.cuda_stream)
        dev = {k: torch.from_numpy(v.view(np.uint8).copy()).cuda() for k, v in cols.items()}
        out = (torch.empty(n, dtype=torch.int32, device="cuda"), torch.empty((n, 2), dtype=torch.int64, device="cuda"),
               torch.zeros(1, dtype=torch.int64, device="cuda"))
        runs.append((cols, stream, clf, dev, out))
    torch.cuda.synchronize()
    for _ in range(3):
        for cols, stream, clf, dev, out in runs:
            clf.classify_device(n, dev["kind"].data_ptr(), dev["reason_id"].data_ptr(), dev["flags"].data_ptr(),
                                dev["cur_stage"].data_ptr(), dev["job_key"].data_ptr(), out[0].data_ptr(),
                                out[1].data_ptr(), out[2].data_ptr())
    torch.cuda.synchronize()
    for cols, stream, clf, dev, out in runs:
        want_v, want_r = cpu_oracle.classify(cols, threads=cpu_oracle.hardware_threads())
        jobs = int(out[2].item())
        got_r = out[1][:jobs].cpu().numpy().view(np.uint64).reshape(-1, 2).copy().view(cpu_oracle.JOB_ROW).reshape(-1)
        util.assert_same(out[0].cpu().numpy().view(np.uint32), got_r, want_v, want_r, "concurrent contexts")
        clf.close()


def test_errors_are_reported_not_thrown(clf):
    with pytest.raises(nx.NxsError) as e:
        clf.classify(util.random_cols(np.random.default_rng(0), (1 << 23) + 1, 2))
    assert e.value.status == nx._abi.NXS_ERR_CAPACITY
    with pytest.raises(nx.NxsError) as e:
        nx.Classifier(device=99)
    assert e.value.status == nx._abi.NXS_ERR_ARG
    # misaligned device column
    import torch
    buf = torch.zeros(1 << 16, dtype=torch.uint8, device="cuda")
    p = buf.data_ptr()
    with pytest.raises(nx.NxsError) as e:
        clf.classify_device(64, p + 1, p + 1024, p + 2048, p + 3072, p + 4096, p + 8192, p + 16384, p + 32768)
    assert e.value.status == nx._abi.NXS_ERR_ARG


# ----------------------------------------------------------------- device-resident path
def _device_columns(torch, cap):
    cap = (cap + 1023) & ~1023
    raw = torch.empty(cap * 13, dtype=torch.uint8, device="cuda")
    base = raw.data_ptr()
    ptr = {"job_key": base, "reason_id": base + cap * 8, "kind": base + cap * 10, "flags": base + cap * 11,
           "cur_stage": base + cap * 12}
    return raw, ptr, cap


def _fetch_columns(torch, raw, cap, n):
    host = raw.cpu().numpy()
    return {"job_key": host[:cap * 8].view(np.uint64)[:n].copy(),
            "reason_id": host[cap * 8:cap * 10].view(np.uint16)[:n].copy(),
            "kind": host[cap * 10:cap * 11][:n].copy(), "flags": host[cap * 11:cap * 12][:n].copy(),
            "cur_stage": host[cap * 12:cap * 13][:n].copy()}


@pytest.mark.parametrize("workload", [nx.WORKLOAD_UNIFORM, nx.WORKLOAD_SKEWED])
def test_device_generator_matches_oracle_stream(clf, workload):
    import torch
    n = 1_000_000
    raw, ptr, cap = _device_columns(torch, n)
    clf.set_stream(torch.cuda.current_stream().cuda_stream)
    kept = clf.generate_device(workload, n, ptr["kind"], ptr["reason_id"], ptr["flags"], ptr["cur_stage"],
                               ptr["job_key"], cap, first=12345)
    torch.cuda.synchronize()
    assert kept == n
    got = _fetch_columns(torch, raw, cap, n)
    want = cpu_oracle.generate(workload, n, first=12345)
    for k in want:
        assert np.array_equal(got[k], want[k]), k


def test_device_generator_shards(clf):
    import torch
    n = 600_000
    raw, ptr, cap = _device_columns(torch, n)
    clf.set_stream(torch.cuda.current_stream().cuda_stream)
    total = 0
    for shard in range(4):
        kept = clf.generate_device(nx.WORKLOAD_SKEWED, n, ptr["kind"], ptr["reason_id"], ptr["flags"],
                                   ptr["cur_stage"], ptr["job_key"], cap, shard=shard, n_shards=4)
        torch.cuda.synchronize()
        got = _fetch_columns(torch, raw, cap, kept)
        want = cpu_oracle.generate(nx.WORKLOAD_SKEWED, n, shard=shard, n_shards=4)
        assert kept == len(want["kind"])
        for k in want:
            assert np.array_equal(got[k], want[k]), (shard, k)
        total += kept
    assert total == n


def _classify_device_full(torch, clf, ptr, n):
    verdict = torch.empty(n, dtype=torch.int32, device="cuda")
    rows = torch.empty((n, 2), dtype=torch.int64, device="cuda")
    n_jobs = torch.zeros(1, dtype=torch.int64, device="cuda")
    clf.classify_device(n, ptr["kind"], ptr["reason_id"], ptr["flags"], ptr["cur_stage"], ptr["job_key"],
                        verdict.data_ptr(), rows.data_ptr(), n_jobs.data_ptr())
    torch.cuda.synchronize()
    return verdict, rows, int(n_jobs.item())


def test_device_path_full_size_bit_exact():
    """BASELINE config 2 at full size (1e8 skewed records, multi-record jobs), device-generated and
    device-resident, compared bit for bit with the oracle in chunks cut on job boundaries; plus the
    size-independent properties: job count == number of key changes, row keys == keys of each job's last
    record, and a second launch reproduces the first."""
    _full_size_check(100_000_000)


@pytest.mark.skipif(os.environ.get("NXS_TEST_1E9", "1") == "0", reason="NXS_TEST_1E9=0")
def test_device_path_1e9_records_bit_exact():
    """The north-star size: 1e9 skewed records in ONE launch on one GPU (33 GB of columns, verdicts and
    rows), every verdict and job row compared with the oracle."""
    import gc
    import torch
    gc.collect()
    torch.cuda.empty_cache()
    free, _ = torch.cuda.mem_get_info()
    if free < 80 * (1 << 30):
        pytest.skip("needs 80 GB of free device memory")
    _full_size_check(1_000_000_000)


@pytest.mark.skipif(os.environ.get("NXS_TEST_MAX", "1") == "0", reason="NXS_TEST_MAX=0")
def test_device_path_maximum_batch_size():
    """The largest batch the ABI takes — 2^32 - 2 records (winner indices are u32 and 0xFFFFFFFF means
    "none") — in ONE launch: 524 288 tiles, 56 GB of columns.  Checked everywhere: the job count and every
    row's key against the keys, every per-record verdict against a table built by the oracle over the whole
    input domain; and bit for bit against the oracle (verdicts and job rows, winner indices included) on
    windows at the start, across record 2^31 and at the very end of the batch."""
    import torch
    n = 2**32 - 2
    import gc
    gc.collect()
    torch.cuda.empty_cache()  # (what earlier tests left in torch's cache counts as used)
    free, _ = torch.cuda.mem_get_info()
    if free < 150 * (1 << 30):
        pytest.skip("needs 150 GB of free device memory")
    step = 1 << 28
    with nx.Classifier(device=0, max_records=n) as clf:
        clf.set_stream(torch.cuda.current_stream().cuda_stream)
        raw, ptr, cap = _device_columns(torch, n)
        assert clf.generate_device(nx.WORKLOAD_SKEWED, n, ptr["kind"], ptr["reason_id"], ptr["flags"],
                                   ptr["cur_stage"], ptr["job_key"], cap) == n
        keys = raw[:cap * 8].view(torch.int64)[:n]

        def tails_of(lo, hi):  # which records of [lo, hi) end a job
            t = torch.ones(hi - lo, dtype=torch.bool, device="cuda")
            last = min(hi, n - 1)
            t[:last - lo] = keys[lo + 1:last + 1] != keys[lo:last]
            return t
        jobs_before = [0]  # job ends before record k * step
        for lo in range(0, n, step):
            jobs_before.append(jobs_before[-1] + int(tails_of(lo, min(n, lo + step)).sum().item()))
        want_jobs = jobs_before[-1]
        verdict = torch.empty(n, dtype=torch.int32, device="cuda")
        rows = torch.empty((want_jobs + 1024, 2), dtype=torch.int64, device="cuda")
        n_jobs = torch.zeros(1, dtype=torch.int64, device="cuda")
        clf.classify_device(n, ptr["kind"], ptr["reason_id"], ptr["flags"], ptr["cur_stage"], ptr["job_key"],
                            verdict.data_ptr(), rows.data_ptr(), n_jobs.data_ptr())
        torch.cuda.synchronize()
        assert clf.last_tile == 8192
        assert int(n_jobs.item()) == want_jobs

        # ---- everywhere: row keys, and the per-record verdicts against the oracle's table of the whole domain
        K, R, F, S = 4, 64, 8, 16
        dom = np.indices((K, R, F, S)).reshape(4, -1)
        table_cols = {"kind": dom[0].astype(np.uint8), "reason_id": dom[1].astype(np.uint16), "flags": dom[2].astype(np.uint8),
                      "cur_stage": dom[3].astype(np.uint8), "job_key": np.arange(dom.shape[1], dtype=np.uint64)}
        table = torch.from_numpy(cpu_oracle.classify(table_cols)[0].view(np.int32)).cuda()
        col = {name: raw[ptr[name] - ptr["job_key"]: ptr[name] - ptr["job_key"] + n * w]
               for name, w in (("reason_id", 2), ("kind", 1), ("flags", 1), ("cur_stage", 1))}
        reason_all = col["reason_id"].view(torch.int16)
        for k, lo in enumerate(range(0, n, step)):
            hi = min(n, lo + step)
            t = tails_of(lo, hi)
            assert torch.equal(rows[jobs_before[k]:jobs_before[k + 1], 0], keys[lo:hi][t]), f"row keys of records [{lo},{hi})"
            del t
            kind, reason = col["kind"][lo:hi].long(), reason_all[lo:hi].long()
            flags, stage = col["flags"][lo:hi].long(), col["cur_stage"][lo:hi].long()
            assert int(kind.max()) < K and int(reason.max()) < R and int(reason.min()) >= 0
            assert int(flags.max()) < F and int(stage.max()) < S
            idx = ((kind * R + reason) * F + flags) * S + stage
            del kind, reason, flags, stage
            assert torch.equal(verdict[lo:hi], table[idx]), f"per-record verdicts of [{lo},{hi})"
            del idx

        # ---- windows against the oracle, verdicts and rows bit for bit
        threads = cpu_oracle.hardware_threads()
        w = 4_000_000
        for start in (0, 2**31 - w // 2, n - w):
            lo = start
            while lo > 0 and int(keys[lo].item()) == int(keys[lo - 1].item()):
                lo += 1  # begin on a job boundary
            hi = min(n, lo + w)
            cols = cpu_oracle.generate(nx.WORKLOAD_SKEWED, hi - lo + 64, first=lo)
            if hi < n:
                k = cols["job_key"]
                while k[hi - lo] == k[hi - lo - 1]:
                    hi += 1
            cols = {name: v[:hi - lo] for name, v in cols.items()}
            assert np.array_equal(keys[lo:hi].cpu().numpy().view(np.uint64), cols["job_key"])
            want_v, want_r = cpu_oracle.classify(cols, threads=threads)
            want_r = want_r.copy()
            has_w = want_r["winner_idx"] != nx.NO_WINNER
            want_r["winner_idx"][has_w] += np.uint32(lo)
            base = lo // step
            row_lo = jobs_before[base] + int(tails_of(base * step, lo).sum().item()) if lo > base * step else jobs_before[base]
            got_v = verdict[lo:hi].cpu().numpy().view(np.uint32)
            got_r = rows[row_lo:row_lo + len(want_r)].cpu().numpy().view(np.uint64).reshape(-1, 2)
            util.assert_same(got_v, got_r.copy().view(cpu_oracle.JOB_ROW).reshape(-1), want_v, want_r,
                             f"records [{lo},{hi})")


def test_one_job_spanning_1e8_records():
    """A batch that is ONE job (every tile but the last is head-less and tail-less: 12 208 tiles whose row
    can only be closed through the look-back over summaries), and the same records cut into three jobs of
    which the middle one spans 12 000 tiles."""
    import torch
    n = 100_000_000
    with nx.Classifier(device=0, max_records=n) as clf:
        clf.set_stream(torch.cuda.current_stream().cuda_stream)
        raw, ptr, cap = _device_columns(torch, n)
        assert clf.generate_device(nx.WORKLOAD_SKEWED, n, ptr["kind"], ptr["reason_id"], ptr["flags"],
                                   ptr["cur_stage"], ptr["job_key"], cap) == n
        cols = cpu_oracle.generate(nx.WORKLOAD_SKEWED, n)
        keys = raw[:cap * 8].view(torch.int64)[:n]
        for cuts in ((), (1_000_003, 99_304_999)):
            key_host = np.full(n, 7, np.uint64)
            for j, c in enumerate(cuts):
                key_host[c:] = 8 + j
            cols["job_key"] = key_host
            keys.copy_(torch.from_numpy(key_host.view(np.int64)))
            verdict, rows, n_jobs = _classify_device_full(torch, clf, ptr, n)
            assert n_jobs == len(cuts) + 1
            want_v, want_r = cpu_oracle.classify(cols, threads=1)
            got_r = rows[:n_jobs].cpu().numpy().view(np.uint64).reshape(-1, 2).copy().view(cpu_oracle.JOB_ROW).reshape(-1)
            util.assert_same(verdict.cpu().numpy().view(np.uint32), got_r, want_v, want_r, f"cuts={cuts}")


def _full_size_check(n):
    import torch
    with nx.Classifier(device=0, max_records=n) as clf:
        clf.set_stream(torch.cuda.current_stream().cuda_stream)
        raw, ptr, cap = _device_columns(torch, n)
        kept = clf.generate_device(nx.WORKLOAD_SKEWED, n, ptr["kind"], ptr["reason_id"], ptr["flags"],
                                   ptr["cur_stage"], ptr["job_key"], cap)
        assert kept == n
        verdict, rows, n_jobs = _classify_device_full(torch, clf, ptr, n)
        keys = raw[:cap * 8].view(torch.int64)[:n]
        tails = torch.ones(n, dtype=torch.bool, device="cuda")
        tails[:-1] = keys[1:] != keys[:-1]
        assert n_jobs == int(tails.sum().item())
        assert torch.equal(rows[:n_jobs, 0], keys[tails])
        del tails
        verdict2, rows2, n_jobs2 = _classify_device_full(torch, clf, ptr, n)
        assert n_jobs2 == n_jobs and torch.equal(verdict2, verdict) and torch.equal(rows2[:n_jobs], rows[:n_jobs])
        del verdict2, rows2

        threads = cpu_oracle.hardware_threads()
        chunk = 10_000_000
        lo, row_lo = 0, 0
        while lo < n:
            hi = min(n, lo + chunk)
            want_cols = cpu_oracle.generate(nx.WORKLOAD_SKEWED, hi - lo + 64, first=lo)
            if hi < n:  # extend to the end of the job that straddles the cut
                k = want_cols["job_key"]
                while k[hi - lo] == k[hi - lo - 1]:
                    hi += 1
            cols = {k: v[:hi - lo] for k, v in want_cols.items()}
            got_keys = keys[lo:hi].cpu().numpy().view(np.uint64)
            assert np.array_equal(got_keys, cols["job_key"]), f"generated keys differ in [{lo},{hi})"
            want_v, want_r = cpu_oracle.classify(cols, threads=threads)
            got_v = verdict[lo:hi].cpu().numpy().view(np.uint32)
            got_r = rows[row_lo:row_lo + len(want_r)].cpu().numpy().view(np.uint64).reshape(-1, 2)
            want_r = want_r.copy()
            has_w = want_r["winner_idx"] != nx.NO_WINNER
            want_r["winner_idx"][has_w] += np.uint32(lo)  # oracle indices are chunk-local
            util.assert_same(got_v, got_r.copy().view(cpu_oracle.JOB_ROW).reshape(-1), want_v, want_r,
                             f"records [{lo},{hi})")
            row_lo += len(want_r)
            lo = hi
        assert row_lo == n_jobs
