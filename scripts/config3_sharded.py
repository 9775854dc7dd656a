#!/usr/bin/env python3
"""BASELINE configs[3]: N records (default 1e9) of the skewed stream sharded by job-id hash across G
GPUs, host-side gather of the verdicts, FULL bit-exact compare against the CPU oracle in chunks.

  python scripts/config3_sharded.py [--records N] [--shards 2,4,8]

The stream is generated once (on GPU 0, 1e8 records at a time, copied to the host), then for every G one
nxs_classify_sharded call partitions it by splitmix64(job_key) % G, classifies each shard on its own GPU
(round-robin when fewer GPUs than shards are present) and merges verdicts and job rows back into stream
order.  The oracle classifies the same host columns chunk by chunk (chunks cut on job boundaries) and
every verdict and every job row is compared.  Needs GPUs; prints one line per G and a JSON summary."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--records", type=int, default=1_000_000_000)
    ap.add_argument("--shards", default="2,4,8")
    ap.add_argument("--chunk", type=int, default=100_000_000)
    args = ap.parse_args()
    import torch
    import nexus_supervisor_b200 as nx
    from oracle import cpu_oracle

    n = args.records
    n_gpus = torch.cuda.device_count()
    threads = cpu_oracle.hardware_threads()
    print(f"{n} records, {n_gpus} GPUs present, {threads} host threads", flush=True)

    # ---- the global stream, device-generated in pieces, kept on the host
    t0 = time.perf_counter()
    cols = {"job_key": np.empty(n, np.uint64), "reason_id": np.empty(n, np.uint16), "kind": np.empty(n, np.uint8),
            "flags": np.empty(n, np.uint8), "cur_stage": np.empty(n, np.uint8)}
    piece = min(n, args.chunk)
    cap = (piece + 1023) & ~1023
    with nx.Classifier(device=0, max_records=cap) as gen:
        gen.set_stream(torch.cuda.current_stream().cuda_stream)
        raw = torch.empty(cap * 13, dtype=torch.uint8, device="cuda")
        b = raw.data_ptr()
        ptr = {"job_key": b, "reason_id": b + cap * 8, "kind": b + cap * 10, "flags": b + cap * 11, "cur_stage": b + cap * 12}
        width = {"job_key": 8, "reason_id": 2, "kind": 1, "flags": 1, "cur_stage": 1}
        for lo in range(0, n, piece):
            m = min(piece, n - lo)
            got = gen.generate_device(nx.WORKLOAD_SKEWED, m, ptr["kind"], ptr["reason_id"], ptr["flags"], ptr["cur_stage"],
                                      ptr["job_key"], cap, first=lo)
            assert got == m
            torch.cuda.synchronize()
            for name, w in width.items():
                host = raw[ptr[name] - b: ptr[name] - b + m * w].cpu().numpy()
                cols[name][lo:lo + m] = host.view(cols[name].dtype)
        del raw
    torch.cuda.empty_cache()
    print(f"stream generated in {time.perf_counter() - t0:.1f} s", flush=True)
    # spot check of the generator against the oracle's own generator
    probe = cpu_oracle.generate(nx.WORKLOAD_SKEWED, 1_000_000, first=max(0, n - 1_000_000))
    for k, v in probe.items():
        assert np.array_equal(v, cols[k][max(0, n - 1_000_000):n]), f"generated {k} differs from the oracle stream"

    verdict = np.empty(n, np.uint32)
    rows = np.empty(n, nx.JOB_ROW)
    results = []
    for G in [int(x) for x in args.shards.split(",")]:
        devices = list(range(min(G, n_gpus)))
        t0 = time.perf_counter()
        with nx.ShardedClassifier(G, devices=devices, max_records=n) as sc:
            t_create = time.perf_counter() - t0
            verdict[:] = 0xDEADBEEF
            times = []
            for rep in range(2):  # the first call also allocates the pinned staging of every shard
                t1 = time.perf_counter()
                v, r = sc.classify(cols, out_verdict=verdict, out_rows=rows)
                times.append(time.perf_counter() - t1)
        n_jobs = len(r)
        # ---- full compare, chunk by chunk (chunks end on job boundaries)
        t2 = time.perf_counter()
        lo, row_lo, bad_v, bad_r = 0, 0, 0, 0
        key = cols["job_key"]
        while lo < n:
            hi = min(n, lo + args.chunk)
            while hi < n and key[hi] == key[hi - 1]:
                hi += 1
            want_v, want_r = cpu_oracle.classify({k: c[lo:hi] for k, c in cols.items()}, threads=threads)
            has = want_r["winner_idx"] != nx.NO_WINNER
            want_r["winner_idx"][has] += np.uint32(lo)  # oracle indices are chunk-local
            bad_v += int(np.count_nonzero(v[lo:hi] != want_v))
            got_r = r[row_lo:row_lo + len(want_r)]
            bad_r += int(np.count_nonzero(got_r != want_r)) if len(got_r) == len(want_r) else len(want_r)
            row_lo += len(want_r)
            lo = hi
        ok = bad_v == 0 and bad_r == 0 and row_lo == n_jobs
        res = {"shards": G, "gpus_used": len(devices), "records": n, "jobs": n_jobs, "ok": ok,
               "verdicts_differing": bad_v, "rows_differing": bad_r, "create_s": round(t_create, 2),
               "first_call_s": round(times[0], 2), "second_call_s": round(times[1], 2),
               "records_per_s_second_call": n / times[1], "compare_s": round(time.perf_counter() - t2, 1)}
        results.append(res)
        print(f"G={G} on {len(devices)} GPUs: {n_jobs} job rows, second call {times[1]:.2f} s = {n / times[1]:.3e} records/s "
              f"(first call {times[0]:.2f} s), bit-exact vs oracle: {ok} ({bad_v} verdicts, {bad_r} rows differ)", flush=True)
    print(json.dumps({"config": "BASELINE configs[3]: records sharded by job-id hash, host-side gather, full compare",
                      "gpus_present": n_gpus, "host_threads": threads, "results": results}))
    return 0 if all(x["ok"] for x in results) else 1


if __name__ == "__main__":
    sys.exit(main())
