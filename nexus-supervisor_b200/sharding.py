"""Host-side sharding for G-GPU runs: records are partitioned by job (never splitting a job) with
shard = splitmix64(job_key) % G, each shard is classified independently, and the per-record verdicts
are scattered back into stream order ("host-side gather"; no collective on the data path).  The unit
of dependence is one job = one checkpoint row (test-resources/checkpoints.cql:22 in the reference;
superviseAction reads and writes only that row, services/supervisor.go:264,:301)."""
import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x):
    """Vectorised splitmix64 over a uint64 array (same function as nxs_shard_of uses)."""
    with np.errstate(over="ignore"):
        z = (np.asarray(x, dtype=np.uint64) + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def shard_ids(job_key, n_shards):
    if n_shards <= 1:
        return np.zeros(len(job_key), dtype=np.uint32)
    return (splitmix64(job_key) % np.uint64(n_shards)).astype(np.uint32)


def partition(cols, n_shards):
    """Returns [(shard_cols, record_index)] per shard; record order inside a shard is stream order."""
    ids = shard_ids(cols["job_key"], n_shards)
    out = []
    for s in range(n_shards):
        idx = np.nonzero(ids == s)[0]
        out.append(({k: np.ascontiguousarray(v[idx]) for k, v in cols.items()}, idx))
    return out


def gather(parts, n_records):
    """Host-side gather.  parts: per shard (verdict, rows, record_index, shard_job_key) where
    record_index maps shard positions to stream positions.  Returns (verdict, rows) as one unsharded run
    over the stream would give them: verdicts in stream order; job rows ordered by the stream position
    of each job's last record, winner_idx translated to stream indices.
    Precondition (same as for a single batch): every job_key forms ONE contiguous run in the stream."""
    verdict = np.zeros(n_records, dtype=np.uint32)
    rows_all, pos_all = [], []
    for v, rows, idx, keys in parts:
        verdict[idx] = v
        if len(idx) == 0:
            continue
        rows = rows.copy()
        has = rows["winner_idx"] != 0xFFFFFFFF
        rows["winner_idx"][has] = idx[rows["winner_idx"][has]].astype(np.uint32)
        tails_local = np.nonzero(np.r_[keys[1:] != keys[:-1], True])[0]
        assert len(tails_local) == len(rows), "shard job rows do not match its key runs"
        rows_all.append(rows)
        pos_all.append(idx[tails_local])
    if not rows_all:
        return verdict, np.zeros(0, dtype=[("job_key", "<u8"), ("verdict", "<u4"), ("winner_idx", "<u4")])
    rows = np.concatenate(rows_all)
    order = np.argsort(np.concatenate(pos_all), kind="stable")
    return verdict, rows[order]
