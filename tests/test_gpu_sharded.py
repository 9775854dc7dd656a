"""GPU parity tests of nxs_classify_sharded (BASELINE config 3): one host batch partitioned by
splitmix64(job_key) % G, every shard classified by the CUDA path on its own ctx, results gathered on the
host — compared bit for bit with ONE unsharded run of the CPU oracle over the whole batch.  G logical
shards map round-robin onto the GPUs present, so every G runs on a single-GPU box too."""
import numpy as np
import pytest

import nexus_supervisor_b200 as nx
from oracle import cpu_oracle
from tests import util

pytestmark = pytest.mark.gpu


def _devices():
    import torch
    return list(range(torch.cuda.device_count()))


@pytest.mark.parametrize("G", [1, 2, 3, 8])
def test_sharded_equals_one_unsharded_oracle_run(G):
    rng = np.random.default_rng(100 + G)
    n = 1_200_003
    with nx.ShardedClassifier(G, devices=_devices(), max_records=n) as sc:
        for max_len, wide in ((6, True), (3000, False)):
            cols = util.random_cols(rng, n, max_len, wide=wide)
            want_v, want_r = cpu_oracle.classify(cols, threads=cpu_oracle.hardware_threads())
            got_v, got_r = sc.classify(cols)
            util.assert_same(got_v, got_r, want_v, want_r, f"G={G} max_len={max_len}")
            none_v, rows_only = sc.classify(cols, want_verdict=False)
            assert none_v is None and np.array_equal(rows_only, want_r)


def test_sharded_synthetic_stream_and_small_batches():
    devs = _devices()
    with nx.ShardedClassifier(4, devices=devs, max_records=3_000_000) as sc:
        for n in (1, 2, 5, 63, 70_001, 3_000_000):
            cols = cpu_oracle.generate(nx.WORKLOAD_SKEWED, n)
            util.assert_same(*sc.classify(cols), *cpu_oracle.classify(cols, threads=4), f"n={n}")
        v, r = sc.classify(util.make_cols([], [], [], [], []))
        assert len(v) == 0 and len(r) == 0


def test_sharded_matches_the_single_context_path_and_shard_function():
    """Same rows as one nxs_classify call, and every job went to the shard nxs_shard_of names."""
    rng = np.random.default_rng(5)
    n = 400_000
    cols = util.random_cols(rng, n, 9)
    with nx.Classifier(device=0, max_records=n) as one, nx.ShardedClassifier(3, devices=_devices(), max_records=n) as sc:
        v1, r1 = one.classify(cols)
        v3, r3 = sc.classify(cols)
        util.assert_same(v3, r3, v1.copy(), r1.copy(), "sharded vs single ctx")
    from nexus_supervisor_b200 import sharding
    ids = sharding.shard_ids(cols["job_key"][:200], 3)
    assert [nx.shard_of(int(k), 3) for k in cols["job_key"][:200]] == ids.tolist()


def test_sharded_errors():
    with pytest.raises(nx.NxsError) as e:
        nx.ShardedClassifier(0)
    assert e.value.status == nx._abi.NXS_ERR_ARG
    with nx.ShardedClassifier(2, devices=_devices(), max_records=1000) as sc:
        with pytest.raises(nx.NxsError) as e:
            sc.classify(util.random_cols(np.random.default_rng(0), 1001, 2))
        assert e.value.status == nx._abi.NXS_ERR_CAPACITY
