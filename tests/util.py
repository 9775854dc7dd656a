"""Shared helpers for the parity tests."""
import numpy as np

from oracle import cpu_oracle

FM_DEFAULT = cpu_oracle.FINISHED_MASK_DEFAULT
MASKS = [FM_DEFAULT, 0, 0xFFFF, 1 << 7, (1 << 7) | (1 << 4), 0x5A5A]


def make_cols(kind, reason, flags, stage, key):
    return {"kind": np.asarray(kind, np.uint8), "reason_id": np.asarray(reason, np.uint16),
            "flags": np.asarray(flags, np.uint8), "cur_stage": np.asarray(stage, np.uint8),
            "job_key": np.asarray(key, np.uint64)}


def random_cols(rng, n, max_job_len, wide=True):
    """Random records; jobs of 1..max_job_len adjacent records."""
    n_jobs = max(1, min(n, 4 * n // (max_job_len + 1) + 16))  # enough jobs to cover n records, bounded memory
    lens = rng.integers(1, max_job_len + 1, size=n_jobs)
    while int(lens.sum()) < n:
        lens = np.concatenate([lens, rng.integers(1, max_job_len + 1, size=n_jobs)])
    lens = lens[:int(np.searchsorted(np.cumsum(lens), n)) + 1]
    with np.errstate(over="ignore"):
        ids = np.arange(1, len(lens) + 1, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15)
    keys = np.repeat(ids, lens)[:n]
    if wide:
        kind = rng.integers(0, 5, n)
        reason = rng.choice([0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 15, 31, 32, 500, 65535], n)
        flags = rng.choice([7, 7, 7, 7, 3, 5, 6, 0, 15, 255], n)
        stage = rng.integers(0, 16, n)
    else:
        kind = rng.integers(1, 3, n)
        reason = rng.integers(1, 9, n)
        flags = np.full(n, 7)
        stage = rng.integers(1, 3, n)
    return make_cols(kind, reason, flags, stage, keys)


def exhaustive_cols(records_per_job=1):
    """Every (kind 0..4) x (reason 0..33 + big) x (flags 0..7) x (stage 0..15) combination."""
    kinds, reasons, flags, stages = np.meshgrid(np.arange(5), np.array(list(range(34)) + [255, 256, 65535]),
                                                np.arange(8), np.arange(16), indexing="ij")
    n = kinds.size
    keys = (np.arange(n, dtype=np.uint64) // np.uint64(records_per_job) + np.uint64(1)) * np.uint64(0x100000001B3)
    return make_cols(kinds.ravel(), reasons.ravel(), flags.ravel(), stages.ravel(), keys)


def assert_same(got_v, got_r, want_v, want_r, what=""):
    assert len(got_v) == len(want_v), what
    bad = np.nonzero(np.asarray(got_v) != np.asarray(want_v))[0]
    assert bad.size == 0, f"{what}: {bad.size} verdicts differ, first at {bad[:5]}: got {got_v[bad[:5]]} want {want_v[bad[:5]]}"
    assert len(got_r) == len(want_r), f"{what}: {len(got_r)} job rows, oracle has {len(want_r)}"
    badr = np.nonzero(got_r != want_r)[0]
    assert badr.size == 0, f"{what}: {badr.size} job rows differ, first at {badr[:5]}: got {got_r[badr[:5]]} want {want_r[badr[:5]]}"
