#!/usr/bin/env python3
"""BASELINE configs[4] — end-to-end replay: classify -> job rows -> checkpoint-row encode -> sink.

The Scylla write path and the kube-apiserver are unavailable on this box (no Docker, no network), so
the sink is a MODEL: a discrete-event simulation of what stands between a verdict and its effect in the
reference — the rate-limited actor of services/supervisor.go:107-117 (token bucket
RateLimitElementsPerSecond / RateLimitElementsBurst, `Workers` goroutines, retry with back-off from
FailureRateBaseDelay doubling up to FailureRateMaxDelay, :41-47; defaults .helm/values.yaml:141-161) in
front of one ReadCheckpoint + one Jobs().Delete + one UpsertCheckpoint per element (:264, :289, :301).

What is NOT modelled but measured: the stream of elements.  It comes from the real pipeline — the
synthetic config-2 stream classified (oracle here, bit-identical to the GPU's rows) and run through the
product's checkpoint-row encoder — so the share of job rows that upsert, delete or fail is computed, not
typed in; classification and encoding rates are read from a bench.py JSON line.

  python scripts/replay_sim.py [bench.json] [--records N]      prints the config-4 table (markdown)
"""
import argparse
import heapq
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


class TokenBucket:
    """`rate` tokens per second, at most `burst` stored; starts full (golang.org/x/time/rate semantics, which the
    nexus-core actor's limiter presumably wraps — unpinned)."""

    def __init__(self, rate, burst):
        self.rate, self.burst = float(rate), float(burst)
        self.tokens, self.t = float(burst), 0.0

    def take(self, now):
        """Earliest time >= now at which one token can be taken; takes it."""
        if self.rate == float("inf"):
            return now
        now = max(now, self.t)
        self.tokens = min(self.burst, self.tokens + (now - self.t) * self.rate)
        self.t = now
        if self.tokens >= 1.0:
            self.tokens -= 1.0
            return now
        wait = (1.0 - self.tokens) / self.rate
        self.t = now + wait
        self.tokens = 0.0
        return self.t


def backoff_delay(attempt, base, cap):
    """Delay before retry number `attempt` (1, 2, ...): base, 2 base, 4 base, ... capped (per-item
    exponential failure rate limiter, client-go workqueue semantics)."""
    return min(cap, base * (2 ** (attempt - 1)))


def simulate(elements, rate=10.0, burst=100, workers=2, latency_s=1e-3, base_delay=0.1, max_delay=1.0, max_retries=3,
             arrival_s=0.0):
    """elements: iterable of booleans — True if the element's processing fails (ReadCheckpoint error: the
    `err` rows, :265-273) and is retried with back-off until max_retries, False if it succeeds at once.
    All elements are queued at `arrival_s` (the GPU hands over a whole batch).  Returns (makespan_s,
    processed_calls, dropped, first_pass_s) — first_pass_s: when every element has had its first attempt
    (the steady-state figure: the last retries of a finite sample trail behind with their back-off)."""
    bucket = TokenBucket(rate, burst)
    free = [arrival_s] * workers  # times at which each worker is free
    heapq.heapify(free)
    queue = [(arrival_s, i, 0, bool(f)) for i, f in enumerate(elements)]  # (ready time, id, attempt, fails)
    heapq.heapify(queue)
    end, calls, dropped, first_pass = arrival_s, 0, 0, arrival_s
    while queue:
        ready, i, attempt, fails = heapq.heappop(queue)
        start = max(ready, heapq.heappop(free))
        start = bucket.take(start)
        done = start + latency_s
        heapq.heappush(free, done)
        calls += 1
        end = max(end, done)
        if attempt == 0:
            first_pass = max(first_pass, done)
        if fails:
            if attempt + 1 > max_retries:
                dropped += 1
            else:
                heapq.heappush(queue, (done + backoff_delay(attempt + 1, base_delay, max_delay), i, attempt + 1, True))
    return end - arrival_s, calls, dropped, first_pass - arrival_s


def pipeline_stream(records):
    """The element stream of `records` config-2 records, from the real pipeline pieces that run without a
    GPU: oracle job rows (bit-identical to the kernel's) -> nxs_encode_checkpoint_rows."""
    import numpy as np
    import nexus_supervisor_b200 as nx
    from oracle import cpu_oracle
    cols = cpu_oracle.generate(nx.WORKLOAD_SKEWED, records)
    _, rows = cpu_oracle.classify(cols, threads=cpu_oracle.hardware_threads())
    v = rows["verdict"]
    acts = v != 0  # rows the actor receives at all
    err = ((v >> 15) & 3) != 0
    arena, offsets, update_row, flags = nx.encode_checkpoint_rows(rows)
    return {"records": int(records), "job_rows": int(len(rows)), "elements": int(acts.sum()),
            "upserts": int(len(update_row)), "deletes": int(((v >> 12) & 1).sum()), "err_rows": int(err.sum()),
            "skips_only": int((acts & (((v >> 13) & 1) == 0) & ~err).sum()),
            "encoded_bytes": int(arena.nbytes), "fails": (err[acts]).astype(bool)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("bench", nargs="?", default=os.path.join(ROOT, "profiles", "r02_final_bench.json"))
    ap.add_argument("--records", type=int, default=10_000_000, help="records of the stream run through the pipeline")
    ap.add_argument("--sim-elements", type=int, default=20_000, help="elements simulated in detail per sink")
    args = ap.parse_args()
    b = json.load(open(args.bench))
    n_bench = b["config"]["records_per_gpu"]
    gpu_rate = b["e2e"]["value"]
    cpu_rate = b.get("cpu_baseline", {}).get("value")
    enc = b.get("checkpoint_row_encoder", {})
    st = pipeline_stream(args.records)
    scale = n_bench / st["records"]
    elements = st["elements"] * scale
    print(f"<!-- python scripts/replay_sim.py {os.path.relpath(args.bench, ROOT)} --records {args.records} -->")
    print(f"Stream (measured on {st['records']:.0e} records of config 2, scaled to {n_bench:.0e}): {st['job_rows'] * scale:.3e} job rows, "
          f"{elements:.3e} reach the actor, {st['upserts'] * scale:.3e} upsert a checkpoint "
          f"({100.0 * st['upserts'] / st['job_rows']:.1f} % of job rows), {st['deletes'] * scale:.3e} delete the job, "
          f"{st['err_rows'] * scale:.3e} have no checkpoint and are retried; the encoder emits "
          f"{st['encoded_bytes'] * scale / 1e9:.2f} GB of column bytes (without the events' messages).")
    t_gpu = n_bench / gpu_rate
    t_cpu = n_bench / cpu_rate if cpu_rate else None
    t_enc = (st["job_rows"] * scale) / enc["value"] if enc.get("value") else 0.0
    print(f"Classification of {n_bench:.0e} records end to end: GPU {t_gpu * 1e3:.1f} ms"
          + (f", CPU event loop {t_cpu:.2f} s" if t_cpu else "") + f"; encoding the rows: {t_enc * 1e3:.0f} ms (host).")
    print()
    print("| sink (MODEL) | elements/s through the sink | time for the batch's elements | records/s, GPU classify | records/s, CPU loop | classification share (GPU) |")
    print("|---|---|---|---|---|---|")
    sinks = [("null sink (upper bound)", None),
             ("reference defaults: 10/s, burst 100, 2 workers, 1 ms per element", dict(rate=10, burst=100, workers=2, latency_s=1e-3)),
             ("rate limit lifted, 2 workers x 0.1 ms", dict(rate=float("inf"), burst=1, workers=2, latency_s=1e-4)),
             ("rate limit lifted, 2 workers x 1 ms", dict(rate=float("inf"), burst=1, workers=2, latency_s=1e-3)),
             ("rate limit lifted, 2 workers x 10 ms", dict(rate=float("inf"), burst=1, workers=2, latency_s=1e-2)),
             ("rate limit lifted, 64 workers x 1 ms", dict(rate=float("inf"), burst=1, workers=64, latency_s=1e-3))]
    k = min(args.sim_elements, len(st["fails"]))
    for name, cfg in sinks:
        if cfg is None:
            t_sink, per_s = 0.0, float("inf")
        else:
            _, calls, dropped, span = simulate(st["fails"][:k], **cfg)
            burst_free = min(cfg["burst"], k) if cfg["rate"] != float("inf") else 0
            # steady-state rate from the simulated sample (the initial burst is a one-off), extrapolated
            per_s = (k - burst_free) / span if span > 0 else float("inf")
            t_sink = max(0.0, elements - burst_free) / per_s
        tot_gpu, tot_cpu = t_gpu + t_enc + t_sink, (t_cpu + t_enc + t_sink) if t_cpu else None
        print(f"| {name} | {'-' if cfg is None else f'{per_s:.3g}'} | {t_sink:.3g} s | {n_bench / tot_gpu:.3g} | "
              f"{(n_bench / tot_cpu) if tot_cpu else float('nan'):.3g} | {100.0 * t_gpu / tot_gpu:.2g} % |")
    print()
    print("Every sink row is a model (the Scylla write path is unavailable on this box); the element stream, the "
          "classification and the encoding times are measured.")


if __name__ == "__main__":
    main()
