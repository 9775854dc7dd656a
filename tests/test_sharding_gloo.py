"""World-size-2 test (gloo, CPU) of the multi-GPU host logic: shard by splitmix64(job_key) % G,
classify each shard independently (here with the oracle — there is no GPU), gather on rank 0 and compare
with one unsharded run.  No collective is involved in the data path itself; gloo only moves the
per-shard results to rank 0, as the host-side gather would."""
import os
import socket

import numpy as np
import pytest

from tests import util


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n, seed, result_path):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    import nexus_supervisor_b200 as nx
    from nexus_supervisor_b200 import sharding
    from oracle import cpu_oracle
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(seed)  # every rank builds the same stream
    cols = util.random_cols(rng, n, 6)
    shard_cols, idx = sharding.partition(cols, world)[rank]
    # the scalar C-ABI function and the vectorised host function must agree
    for k in shard_cols["job_key"][:50]:
        assert nx.shard_of(int(k), world) == rank
    v, rows = cpu_oracle.classify(shard_cols)
    gathered = [None] * world if rank == 0 else None
    dist.gather_object((v, rows, idx, shard_cols["job_key"]), gathered, dst=0)
    if rank == 0:
        got_v, got_r = sharding.gather(gathered, n)
        want_v, want_r = cpu_oracle.classify(cols)
        ok = np.array_equal(got_v, want_v) and np.array_equal(got_r, want_r)
        np.save(result_path, np.array([int(ok), len(want_r), sum(len(g[2]) for g in gathered)]))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2])
def test_sharded_equals_unsharded(tmp_path, world):
    import torch.multiprocessing as mp
    n = 50_000
    result = str(tmp_path / "result.npy")
    mp.spawn(_worker, args=(world, _free_port(), n, 11, result), nprocs=world, join=True)
    ok, n_rows, n_seen = np.load(result)
    assert ok == 1 and n_seen == n and n_rows > 0


def test_partition_keeps_jobs_whole_and_ordered():
    from nexus_supervisor_b200 import sharding
    from oracle import cpu_oracle
    cols = cpu_oracle.generate(2, 100_000)
    parts = sharding.partition(cols, 8)
    assert sum(len(idx) for _, idx in parts) == 100_000
    seen = set()
    for s, (c, idx) in enumerate(parts):
        assert (np.diff(idx) > 0).all()
        keys = set(np.unique(c["job_key"]).tolist())
        assert not (keys & seen)
        seen |= keys
        want = cpu_oracle.generate(2, 100_000, shard=s, n_shards=8)
        for k in want:
            assert np.array_equal(c[k], want[k])
    sizes = np.array([len(idx) for _, idx in parts])
    assert sizes.max() / sizes.mean() < 1.05


def test_division_free_shard_map_equals_shard_of():
    """The host passes of nxs_classify_sharded compute splitmix64(key) % G without a division (mask, or a
    folded remainder by multiplication): same value as nxs_shard_of for every G the entry point accepts."""
    import ctypes as C
    import nexus_supervisor_b200 as nx
    lib = nx._abi.load()
    lib.nxs_internal_shard_map.restype = None
    lib.nxs_internal_shard_map.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p]
    rng = np.random.default_rng(5)
    keys = rng.integers(0, 2**64, 4096, dtype=np.uint64)
    keys[:8] = [0, 1, 2**32 - 1, 2**32, 2**63, 2**64 - 1, 0x6E65787573, 12345]
    got = np.empty(len(keys), np.uint32)
    for G in range(1, 257):
        lib.nxs_internal_shard_map(keys.ctypes.data, len(keys), G, got.ctypes.data)
        want = np.array([nx.shard_of(int(k), G) for k in keys[:64]], np.uint32)
        assert np.array_equal(got[:64], want), G
        # and the whole array against the definition, vectorised (splitmix64 in numpy)
        with np.errstate(over="ignore"):
            z = keys + np.uint64(0x9E3779B97F4A7C15)
            z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
            z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
            z = z ^ (z >> np.uint64(31))
        assert np.array_equal(got, (z % np.uint64(G)).astype(np.uint32)), G


@pytest.mark.parametrize("G", [1, 2, 3, 8])
def test_product_partition_and_gather_with_the_oracle_as_classifier(G):
    """The host logic of nxs_classify_sharded itself (C++: parallel partition on job boundaries, per-shard
    classification, verdict scatter, job-row merge with winner indices translated back), run on the CPU
    through the library's test hook with the oracle standing in for the kernel, against one unsharded
    oracle run."""
    import ctypes as C
    import nexus_supervisor_b200 as nx
    from oracle import cpu_oracle
    lib = nx._abi.load()
    FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_uint32, C.POINTER(nx._abi.Batch), C.POINTER(nx._abi.Out))
    lib.nxs_internal_sharded_run.restype = C.c_int
    lib.nxs_internal_sharded_run.argtypes = [C.c_uint32, C.POINTER(nx._abi.Batch), C.POINTER(nx._abi.Out), FN, C.c_void_p]
    seen = {}

    def classify_shard(user, shard, bp, op):
        b, o = bp.contents, op.contents
        n = b.n

        def arr(ptr, dt):
            return np.frombuffer((C.c_uint8 * (n * np.dtype(dt).itemsize)).from_address(ptr), dtype=dt)
        cols = {"kind": arr(b.kind, np.uint8), "reason_id": arr(b.reason_id, np.uint16), "flags": arr(b.flags, np.uint8),
                "cur_stage": arr(b.cur_stage, np.uint8), "job_key": arr(b.job_key, np.uint64)}
        assert all(nx.shard_of(int(k), G) == shard for k in cols["job_key"][:20])
        v, r = cpu_oracle.classify(cols)
        if o.verdict:
            arr(o.verdict, np.uint32)[:] = v
        np.frombuffer((C.c_uint8 * (16 * len(r))).from_address(o.job_rows), dtype=cpu_oracle.JOB_ROW)[:] = r
        C.c_uint64.from_address(o.n_jobs).value = len(r)
        seen[shard] = n
        return 0

    rng = np.random.default_rng(G)
    cases = [util.random_cols(rng, n, max_len) for n, max_len in ((1, 1), (777, 3), (40_000, 6), (30_001, 900))]
    one_job = util.random_cols(rng, 5_000, 1, wide=False)
    one_job["job_key"][:] = 99  # a single job: one shard gets everything, the others nothing
    few = util.random_cols(rng, 3, 1)  # fewer records than shards
    cases += [one_job, few]
    for cols in cases:
        n = len(cols["kind"])
        verdict = np.zeros(n, np.uint32)
        rows = np.zeros(n, cpu_oracle.JOB_ROW)
        nj = C.c_uint64(0)
        batch = nx._abi.Batch(cols["kind"].ctypes.data, cols["reason_id"].ctypes.data, cols["flags"].ctypes.data,
                              cols["cur_stage"].ctypes.data, cols["job_key"].ctypes.data, n)
        out = nx._abi.Out(verdict.ctypes.data, rows.ctypes.data, C.addressof(nj))
        seen.clear()
        assert lib.nxs_internal_sharded_run(G, C.byref(batch), C.byref(out), FN(classify_shard), None) == 0
        assert sum(seen.values()) == n
        util.assert_same(verdict, rows[:nj.value], *cpu_oracle.classify(cols), f"G={G} n={n}")
        # job rows only (out->verdict == NULL): nothing per record is asked of the shards or merged
        rows2 = np.zeros(n, cpu_oracle.JOB_ROW)
        out2 = nx._abi.Out(None, rows2.ctypes.data, C.addressof(nj))
        assert lib.nxs_internal_sharded_run(G, C.byref(batch), C.byref(out2), FN(classify_shard), None) == 0
        assert np.array_equal(rows2[:nj.value], rows[:nj.value])
