"""The batched checkpoint-row encoder (nxs_encode_checkpoint_rows): the bytes it puts in
its arena for every upserting job row — lifecycle_stage, algorithm_failure_cause,
algorithm_failure_details — against what the string-mode oracle's replay of the same events leaves in
the checkpoint rows.  Host code on both sides: runs without a GPU (job rows come from the oracle)."""
import numpy as np
import pytest

import nexus_supervisor_b200 as nx
from oracle import cpu_oracle
from tests import util


def messages(n):
    """event.Message of record i = "m<i>", as a heap + offsets."""
    texts = [b"m%d" % i for i in range(n)]
    lens = np.fromiter((len(t) for t in texts), dtype=np.uint64, count=n)
    offs = np.zeros(n + 1, np.uint64)
    np.cumsum(lens, out=offs[1:])
    return np.frombuffer(b"".join(texts), dtype=np.uint8), offs


def check(cols, fm):
    n = len(cols["kind"])
    _, rows = cpu_oracle.classify(cols, fm, threads=4)
    heap, offs = messages(n)
    arena, offsets, update_row, flags = nx.encode_checkpoint_rows(rows, heap, offs)
    want_arena, want_offsets, want_jobs = cpu_oracle.encode_strings(cols, fm)
    assert len(update_row) == len(want_jobs)
    assert np.array_equal(update_row, want_jobs)  # the same jobs were written, in the same order
    assert np.array_equal(offsets, want_offsets)
    assert np.array_equal(arena, want_arena)  # byte for byte
    ups = (rows["verdict"] >> 13) & 1
    assert len(update_row) == int(ups.sum())
    cause = (rows["verdict"][update_row] >> 7) & 3
    assert np.array_equal(flags & 1, (cause != 0).astype(np.uint8))
    return arena, offsets, update_row


def test_encoder_million_rows_byte_exact():
    cols = cpu_oracle.generate(nx.WORKLOAD_SKEWED, 2_500_000)  # ~1.04M job rows, ~0.63M of them upsert
    arena, offsets, update_row = check(cols, util.FM_DEFAULT)
    assert len(update_row) > 500_000
    # spot-check one failure row in full: the reference's doubled sentence for Job / PodFailurePolicy
    text = arena.tobytes()
    assert (b"FAILEDAlgorithm encountered a fatal error during execution: Algorithm encountered a fatal error "
            b"during execution.m") in text


@pytest.mark.parametrize("fm", [0, 0xFFFF, 0x5A5A])
def test_encoder_wide_domain_and_masks(fm):
    """Unusual finished-masks: a failure followed by an applied ToRunning keeps the failure's cause and
    details under the RUNNING stage, as the reference's row would."""
    rng = np.random.default_rng(fm)
    for max_len in (1, 4, 40):
        cols = util.random_cols(rng, 200_000, max_len, wide=True)
        cols["cur_stage"] &= 7  # stages 8..15 have no spelling in the library's table
        keys = cols["job_key"]  # one checkpoint row per job: its records see one stage and one "exists"
        head = np.r_[True, keys[1:] != keys[:-1]]
        first = np.maximum.accumulate(np.where(head, np.arange(len(keys)), 0))
        cols["cur_stage"] = cols["cur_stage"][first]
        cols["flags"] = (cols["flags"] & 3) | (cols["flags"][first] & 4)
        check(cols, fm)


def test_encoder_size_query_and_capacity():
    import ctypes as C
    cols = cpu_oracle.generate(nx.WORKLOAD_SKEWED, 10_000)
    _, rows = cpu_oracle.classify(cols)
    lib = nx._abi.load()
    n_up, n_bytes = C.c_uint64(0), C.c_uint64(0)
    rc = lib.nxs_encode_checkpoint_rows(rows.ctypes.data, len(rows), None, None, 0, None, None, 0, None, None, None,
                                        C.byref(n_up), C.byref(n_bytes))
    assert rc == 0 and n_up.value > 0 and n_bytes.value > 0
    small = np.empty(8, np.uint8)
    offs = np.empty(3 * n_up.value + 1, np.uint64)
    upd = np.empty(n_up.value, np.uint32)
    rc = lib.nxs_encode_checkpoint_rows(rows.ctypes.data, len(rows), None, None, 0, None, small.ctypes.data, 8,
                                        offs.ctypes.data, upd.ctypes.data, None, C.byref(n_up), C.byref(n_bytes))
    assert rc == nx._abi.NXS_ERR_CAPACITY


def test_encoder_into_kept_buffers():
    """out=: a caller that encodes batch after batch writes into buffers it keeps; same bytes as a fresh call,
    stale content of an earlier, larger batch does not leak, a short arena is refused."""
    big = cpu_oracle.generate(nx.WORKLOAD_SKEWED, 200_000)
    small = cpu_oracle.generate(nx.WORKLOAD_SKEWED, 30_000, first=1_000_000)
    bufs = None
    for cols in (big, small):
        _, rows = cpu_oracle.classify(cols)
        n = len(cols["kind"])
        heap = np.frombuffer(b"".join(b"m%d" % i for i in range(n)), np.uint8)
        lens = np.array([len(b"m%d" % i) for i in range(n)], np.uint64)
        offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
        want = nx.encode_checkpoint_rows(rows, heap, offs)
        if bufs is None:
            m = len(rows)
            bufs = (np.full(want[0].nbytes + 100, 0xEE, np.uint8), np.full(3 * m + 1, 7, np.uint64),
                    np.full(m, 9, np.uint32), np.full(m, 3, np.uint8))
        got = nx.encode_checkpoint_rows(rows, heap, offs, out=bufs)
        assert all(np.array_equal(a, b) for a, b in zip(got, want))
    with pytest.raises(nx.NxsError) as e:
        _, rows = cpu_oracle.classify(big)
        nx.encode_checkpoint_rows(rows, out=(np.empty(16, np.uint8),) + bufs[1:])
    assert e.value.status == nx._abi.NXS_ERR_CAPACITY
