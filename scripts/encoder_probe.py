#!/usr/bin/env python3
"""Where the time of nxs_encode_checkpoint_rows goes: the full call, the call without the events' messages
(no details column: no data-dependent reads), and a batch in which no row upserts (the bare walk)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import nexus_supervisor_b200 as nx  # noqa: E402
from oracle import cpu_oracle  # noqa: E402

n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 24_000_000
cols = cpu_oracle.generate(nx.WORKLOAD_SKEWED, n)
_, rows = cpu_oracle.classify(cols, threads=cpu_oracle.hardware_threads())
m = len(rows)
heap = np.full(n * 24, 97, np.uint8)
offs = np.arange(n + 1, dtype=np.uint64) * np.uint64(24)
first = nx.encode_checkpoint_rows(rows, heap, offs)
bufs = (np.zeros(first[0].nbytes + 4096, np.uint8), np.zeros(3 * m + 1, np.uint64), np.zeros(m, np.uint32), np.zeros(m, np.uint8))
none = rows.copy()
none["verdict"] = 0
for name, r, h, o in (("full", rows, heap, offs), ("no messages", rows, None, None), ("no row upserts", none, heap, offs)):
    best = 1e9
    for _ in range(4):
        t0 = time.perf_counter()
        out = nx.encode_checkpoint_rows(r, h, o, out=bufs)
        best = min(best, time.perf_counter() - t0)
    print(f"{name:15s}: {m} rows, {len(out[2])} updates, {out[0].nbytes / 1e6:.0f} MB: {best * 1e3:.1f} ms = {m / best:.3e} rows/s", flush=True)
