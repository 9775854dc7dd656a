#!/usr/bin/env python3
"""nxs_pack_columns on this machine's host cores (records/s)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import nexus_supervisor_b200 as nx  # noqa: E402
from oracle import cpu_oracle  # noqa: E402

n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000_000
cols = cpu_oracle.generate(nx.WORKLOAD_SKEWED, n)
out = np.zeros(n, np.uint16)
for _ in range(4):
    t = time.perf_counter()
    nx.pack_columns(cols, out=out)
    dt = time.perf_counter() - t
    print(f"pack {n} records: {dt * 1e3:.1f} ms = {n / dt:.3e} records/s", flush=True)
