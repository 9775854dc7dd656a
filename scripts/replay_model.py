#!/usr/bin/env python3
"""End-to-end replay MODEL (BASELINE.json configs[4]): classify -> checkpoint-row encode -> sink.

The reference's sink is one kube-apiserver DELETE and one CQL upsert per run behind a token bucket
(10 elements/s, burst 100, 2 workers by default: .helm/values.yaml:141-161,
services/supervisor.go:41-47,107-117).  Running Scylla or the Go supervisor is unavailable on this box, so
the sink is MODELLED: a null sink (upper bound), the reference's own default rate limit, and an assumed
per-upsert latency of 0.1 / 1 / 10 ms on 2 workers.  The point of the table: whatever the classification
step costs, the supervisor stays I/O-bound by 5 to 8 orders of magnitude.

  python scripts/replay_model.py [profiles/r01_v3_bench.json]
"""
import json
import sys


def main():
    path = sys.argv[1] if len(sys.argv) > 1 else "profiles/r01_v3_bench.json"
    b = json.load(open(path))
    n = b["config"]["records_per_gpu"]
    jobs = b["config"]["jobs_per_gpu"]
    # share of jobs whose row calls for an upsert in the skewed stream (oracle, 2e6 records: 60.5 %;
    # 20.8 % also delete the Kubernetes job)
    upsert_share = 0.605
    writes = jobs * upsert_share
    t_gpu = n / b["e2e"]["value"]                 # host buffers in, verdicts + rows out
    t_cpu = n / b["cpu_baseline"]["value"]        # string-faithful C++ stand-in for the Go loop, all cores
    rows = []
    for name, t_sink in (("null sink (upper bound)", 0.0),
                         ("reference default: token bucket 10 rows/s, burst 100", max(0.0, (writes - 100) / 10.0)),
                         ("2 workers x 0.1 ms per upsert (assumed)", writes * 1e-4 / 2),
                         ("2 workers x 1 ms per upsert (assumed)", writes * 1e-3 / 2),
                         ("2 workers x 10 ms per upsert (assumed)", writes * 1e-2 / 2)):
        rows.append({"sink": name, "sink_s": t_sink, "gpu_total_s": t_gpu + t_sink, "cpu_total_s": t_cpu + t_sink,
                     "gpu_records_per_s": n / (t_gpu + t_sink), "cpu_records_per_s": n / (t_cpu + t_sink),
                     "classification_share_gpu": t_gpu / (t_gpu + t_sink)})
    out = {"model": True, "records": n, "jobs": jobs, "writes": int(writes), "classify_e2e_gpu_s": t_gpu,
           "classify_cpu_standin_s": t_cpu, "rows": rows}
    print(json.dumps(out, indent=1))
    print("\n| sink (MODEL) | sink time | records/s with GPU classify | records/s with CPU stand-in | classification share (GPU) |")
    print("|---|---|---|---|---|")
    for r in rows:
        print(f"| {r['sink']} | {r['sink_s']:.3g} s | {r['gpu_records_per_s']:.3g} | {r['cpu_records_per_s']:.3g} | "
              f"{100 * r['classification_share_gpu']:.2g} % |")


if __name__ == "__main__":
    main()
