"""The host flattening layer at scale: a million Kubernetes-Event-shaped structs of
strings through the C++ mirror of the service interface — Supervisor::OnEvent (guards' lookups,
interning, run grouping) and Flush (ReadCheckpoint per run, nxs_classify on the GPU, Delete / Upsert per
job row) — and every applied verdict compared with the string-mode oracle fed the same events."""
import json
import os
import subprocess

import numpy as np
import pytest

import nexus_supervisor_b200 as nx
from oracle import cpu_oracle

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "nexus-supervisor_b200", "lib", "flatten_bench")


def write_columns(path, cols):
    with open(path, "wb") as f:
        f.write(np.uint64(len(cols["kind"])).tobytes())
        for name in ("kind", "reason_id", "flags", "cur_stage", "job_key"):
            f.write(np.ascontiguousarray(cols[name]).tobytes())


def run_flatten(tmp_path, cols, interleave):
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "nexus-supervisor_b200", "host")])
    src, dst = str(tmp_path / "cols.bin"), str(tmp_path / "rows.bin")
    write_columns(src, cols)
    out = subprocess.run([BIN, src, dst] + (["interleave"] if interleave else []), capture_output=True, text=True,
                         timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    stats = json.loads(out.stdout.strip().splitlines()[-1])
    got = np.fromfile(dst, dtype=[("key", "<u8"), ("verdict", "<u4"), ("winner", "<u4")])
    return stats, got


def expected_rows(cols):
    """String-mode oracle over the same records.  Events whose involved object is missing from the cache
    (stale events) belong to no run on the host side; the oracle drops them too, inside their job."""
    v, rows = cpu_oracle.classify(cols, mode="strings", threads=cpu_oracle.hardware_threads())
    rows = rows[rows["verdict"] != 0]
    return np.sort(rows, order="job_key")


@pytest.mark.parametrize("interleave", [False, True])
def test_million_events_through_the_host_mirror(tmp_path, interleave):
    n = 1_000_000
    cols = cpu_oracle.generate(nx.WORKLOAD_SKEWED, n)
    stats, got = run_flatten(tmp_path, cols, interleave)
    assert stats["events"] == n
    want = expected_rows(cols)
    got = np.sort(got, order="key")
    assert len(got) == len(want), (len(got), len(want))
    assert np.array_equal(got["key"], want["job_key"])
    assert np.array_equal(got["verdict"], want["verdict"])
    # the winner: the oracle names a stream index, the mirror resolved it to that event's own message
    has = want["winner_idx"] != nx.NO_WINNER
    assert np.array_equal(got["winner"][has], want["winner_idx"][has])
    assert (got["winner"][~has] == nx.NO_WINNER).all()
    print("flatten stats:", stats)
