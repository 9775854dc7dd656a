"""GPU parity tests of the packed batch type (job_key u64 + attr u16) and of the rows-only mode
(out->verdict == NULL): both exist to cut the bytes that cross PCIe, and both must give exactly what the
oracle gives for the unpacked columns."""
import numpy as np
import pytest

import nexus_supervisor_b200 as nx
from oracle import cpu_oracle
from tests import util

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("fm", [util.FM_DEFAULT, 0, 0x5A5A])
def test_packed_exhaustive_domain(fm):
    """Every (kind 0..4, reason 0..33 / 255 / 256 / 65535, flags 0..7, stage 0..15): packing clamps kinds >= 3
    and reasons >= 63, which the reference's switches cannot tell from the unclamped values."""
    with nx.Classifier(device=0, max_records=1 << 18, finished_mask=fm) as c:
        for per_job in (1, 3):
            cols = util.exhaustive_cols(per_job)
            want_v, want_r = cpu_oracle.classify(cols, fm)
            got_v, got_r = c.classify_packed(cols["job_key"], nx.pack_columns(cols), want_verdict=True)
            util.assert_same(got_v, got_r, want_v, want_r, f"packed fm={fm:#x} per_job={per_job}")
            none_v, rows_only = c.classify_packed(cols["job_key"], nx.pack_columns(cols))
            assert none_v is None and np.array_equal(rows_only, want_r)


def test_packed_and_rows_only_sliced_batches():
    """20M records: several 8M-record slices through the packed upload and the attr expansion."""
    rng = np.random.default_rng(41)
    n = 20_000_003
    cols = util.random_cols(rng, n, 7)
    want_v, want_r = cpu_oracle.classify(cols, threads=cpu_oracle.hardware_threads())
    attr = nx.pack_columns(cols)
    assert np.array_equal(attr & 63, np.minimum(cols["reason_id"], 63)) and np.array_equal((attr >> 11) & 15, cols["cur_stage"] & 15)
    with nx.Classifier(device=0, max_records=n) as c:
        v, r = c.classify_packed(cols["job_key"], attr, want_verdict=True)
        util.assert_same(v, r, want_v, want_r, "packed, sliced")
        _, r2 = c.classify_packed(cols["job_key"], attr)
        assert np.array_equal(r2, want_r)
        _, r3 = c.classify(cols, want_verdict=False)  # unpacked columns, rows only
        assert np.array_equal(r3, want_r)
        v4, r4 = c.classify(cols)  # and the ordinary call still returns both
        util.assert_same(v4, r4, want_v, want_r, "after rows-only calls")


def test_rows_only_device_path():
    import torch
    rng = np.random.default_rng(43)
    n = 3_000_001
    cols = util.random_cols(rng, n, 30)
    want_v, want_r = cpu_oracle.classify(cols, threads=4)
    with nx.Classifier(device=0, max_records=n) as c:
        dev = {k: torch.from_numpy(v.view(np.uint8).copy()).cuda() for k, v in cols.items()}
        rows = torch.empty((n, 2), dtype=torch.int64, device="cuda")
        nj = torch.zeros(1, dtype=torch.int64, device="cuda")
        c.classify_device(n, dev["kind"].data_ptr(), dev["reason_id"].data_ptr(), dev["flags"].data_ptr(),
                          dev["cur_stage"].data_ptr(), dev["job_key"].data_ptr(), 0, rows.data_ptr(), nj.data_ptr())
        c.sync()
        jobs = int(nj.item())
        got = rows[:jobs].cpu().numpy().view(np.uint64).reshape(-1, 2).copy().view(cpu_oracle.JOB_ROW).reshape(-1)
        assert np.array_equal(got, want_r)


def test_page_locked_caller_buffers_sliced():
    """The fast host path: the caller's columns and outputs are page-locked (torch pinned memory here, what
    nxs_alloc_host hands out), so the library copies straight from and to them, slice by slice, with no
    bounce buffers in between — default call, rows only, and packed."""
    import torch
    rng = np.random.default_rng(47)
    n = 19_000_017
    cols = util.random_cols(rng, n, 9)
    want_v, want_r = cpu_oracle.classify(cols, threads=cpu_oracle.hardware_threads())

    def pin(a):
        t = torch.from_numpy(a.view(np.uint8).reshape(-1).copy()).pin_memory()
        assert t.is_pinned()
        return t.numpy().view(a.dtype).reshape(a.shape), t  # (the tensor keeps the memory alive)
    held = []
    pinned = {}
    for k, v in cols.items():
        pinned[k], t = pin(v)
        held.append(t)
    out_v, tv = pin(np.full(n, 0xDEADBEEF, np.uint32))
    out_r, tr = pin(np.zeros(n, cpu_oracle.JOB_ROW))
    attr, ta = pin(nx.pack_columns(cols))
    held += [tv, tr, ta]
    with nx.Classifier(device=0, max_records=n) as c:
        v, r = c.classify(pinned, out_verdict=out_v, out_rows=out_r)
        util.assert_same(v, r, want_v, want_r, "page-locked buffers, sliced")
        out_r[:] = np.zeros(1, cpu_oracle.JOB_ROW)
        _, r2 = c.classify(pinned, out_rows=out_r, want_verdict=False)
        assert np.array_equal(r2, want_r)
        out_r[:] = np.zeros(1, cpu_oracle.JOB_ROW)
        out_v[:] = 0
        v3, r3 = c.classify_packed(pinned["job_key"], attr, out_verdict=out_v, out_rows=out_r, want_verdict=True)
        util.assert_same(v3, r3, want_v, want_r, "page-locked buffers, packed")


@pytest.mark.parametrize("fm", [util.FM_DEFAULT, 0, 0x5A5A])
def test_multi_slice_batch_over_the_whole_domain(fm):
    """Batches of more than one slice take the link-lean route inside nxs_classify — the four small columns
    travel packed, the per-record verdicts come back as one byte each and are expanded on the host — which must
    not change a bit of the result: every (kind, reason, flags, stage) combination, repeated until the batch
    needs two slices, jobs of 1 and of 3 records, three finished-masks."""
    for per_job in (1, 3):
        base = util.exhaustive_cols(per_job)
        reps = (9_437_184 + 8192) // len(base["kind"]) + 1
        cols = {k: np.tile(v, reps) for k, v in base.items()}
        n = len(cols["kind"])
        with np.errstate(over="ignore"):  # distinct keys per repetition, runs stay adjacent
            cols["job_key"] = cols["job_key"] + np.repeat(np.arange(reps, dtype=np.uint64) << np.uint64(40), len(base["kind"]))
        want_v, want_r = cpu_oracle.classify(cols, fm, threads=cpu_oracle.hardware_threads())
        with nx.Classifier(device=0, max_records=n, finished_mask=fm) as c:
            got_v, got_r = c.classify(cols)
            util.assert_same(got_v, got_r, want_v, want_r, f"two slices fm={fm:#x} per_job={per_job}")
            up, down = c.last_wire_bytes
            assert up == 10 * n and down < 2 * n + 16 * len(want_r)
