#!/usr/bin/env python3
"""What limits the end-to-end (host buffers) numbers when several GPUs share one host: host DRAM
bandwidth, or the PCIe path?  Measures, on this box:

  1. plain host memcpy bandwidth, 1 thread and all threads (numpy copies release the GIL);
  2. page-locked host <-> device copy bandwidth of every GPU ALONE (H2D, D2H, both at once);
  3. the same for ALL GPUs at the same time (aggregate and per GPU).

  python scripts/host_link_bw.py [--mb 1024]      prints one JSON object.  Needs GPUs.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np


def host_memcpy(mb, threads):
    n = mb << 20
    src = [np.ones(n, np.uint8) for _ in range(threads)]
    dst = [np.empty(n, np.uint8) for _ in range(threads)]
    best = 0.0
    for _ in range(3):
        ts = [threading.Thread(target=np.copyto, args=(dst[i], src[i])) for i in range(threads)]
        t0 = time.perf_counter()
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        dt = time.perf_counter() - t0
        best = max(best, threads * n / dt / 1e9)
    return best  # GB/s copied (read + write traffic is twice that)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=1024)
    args = ap.parse_args()
    import torch
    n_gpus = torch.cuda.device_count()
    cores = os.cpu_count() or 1
    out = {"gpus": n_gpus, "host_threads": cores, "buffer_mb": args.mb,
           "host_memcpy_GBps": {"1_thread": host_memcpy(256, 1), f"{min(cores, 16)}_threads": host_memcpy(256, min(cores, 16))}}
    n = args.mb << 20
    host_in = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(n_gpus)]
    host_out = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(n_gpus)]
    dev_a = [torch.empty(n, dtype=torch.uint8, device=f"cuda:{g}") for g in range(n_gpus)]
    dev_b = [torch.empty(n, dtype=torch.uint8, device=f"cuda:{g}") for g in range(n_gpus)]
    s_up = [torch.cuda.Stream(device=g) for g in range(n_gpus)]
    s_dn = [torch.cuda.Stream(device=g) for g in range(n_gpus)]

    def run(gpus, up, down, reps=3):
        best = None
        for _ in range(reps + 1):
            for g in gpus:
                torch.cuda.synchronize(g)
            t0 = time.perf_counter()
            for g in gpus:
                if up:
                    with torch.cuda.stream(s_up[g]):
                        dev_a[g].copy_(host_in[g], non_blocking=True)
                if down:
                    with torch.cuda.stream(s_dn[g]):
                        host_out[g].copy_(dev_b[g], non_blocking=True)
            for g in gpus:
                torch.cuda.synchronize(g)
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        return len(gpus) * n / best / 1e9  # GB/s per direction, all listed GPUs together

    alone = {}
    for g in range(n_gpus):
        alone[f"gpu{g}"] = {"h2d": run([g], True, False), "d2h": run([g], False, True), "both_each_way": run([g], True, True)}
    out["one_gpu_at_a_time_GBps"] = alone
    if n_gpus > 1:
        for k in sorted({2, 4, n_gpus} & set(range(2, n_gpus + 1))):
            gs = list(range(k))
            out[f"{k}_gpus_together_GBps"] = {"h2d_total": run(gs, True, False), "d2h_total": run(gs, False, True),
                                              "both_each_way_total": run(gs, True, True)}
    print(json.dumps(out))


if __name__ == "__main__":
    sys.exit(main())
