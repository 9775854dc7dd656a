#!/usr/bin/env python3
"""BASELINE configs[1]: 1M records of the uniform workload (one record per job), classification only on one
B200.  208 launches rotating over 16 distinct buffer sets (16 x 33 MB > L2), CUDA events; plus single cold
launches (L2 flushed by writing 512 MB before each).  Needs a GPU."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import nexus_supervisor_b200 as nx  # noqa: E402


def main():
    n, sets = 1_000_000, 16
    cap = (n + 1023) & ~1023
    clf = nx.Classifier(device=0, max_records=cap)
    stream = torch.cuda.current_stream()
    clf.set_stream(stream.cuda_stream)
    bufs = []
    for k in range(sets):
        raw = torch.empty(cap * 13, dtype=torch.uint8, device="cuda")
        b = raw.data_ptr()
        ptr = dict(job_key=b, reason_id=b + cap * 8, kind=b + cap * 10, flags=b + cap * 11, cur_stage=b + cap * 12)
        clf.generate_device(nx.WORKLOAD_UNIFORM, n, ptr["kind"], ptr["reason_id"], ptr["flags"], ptr["cur_stage"],
                            ptr["job_key"], cap, first=k * n)
        verdict = torch.empty(n, dtype=torch.int32, device="cuda")
        rows = torch.empty((n, 2), dtype=torch.int64, device="cuda")
        bufs.append((raw, ptr, verdict, rows))
    nj = torch.zeros(1, dtype=torch.int64, device="cuda")

    def step(k):
        _, ptr, verdict, rows = bufs[k % sets]
        clf.classify_device(n, ptr["kind"], ptr["reason_id"], ptr["flags"], ptr["cur_stage"], ptr["job_key"],
                            verdict.data_ptr(), rows.data_ptr(), nj.data_ptr())

    for k in range(32):
        step(k)
    torch.cuda.synchronize()
    reps = 208
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for k in range(reps):
        step(k)
    e1.record(stream)
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / reps * 1e3
    bytes_per = 17 * n + 16 * int(nj.item())
    print(f"CONFIG1 rotating: {us:.1f} us per launch, {n / us * 1e6:.3e} records/s, {bytes_per / us * 1e-3:.0f} GB/s "
          f"algorithmic ({int(nj.item())} jobs)")
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    cold = []
    for k in range(8):
        flush.fill_(k)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        step(k)
        b.record(stream)
        torch.cuda.synchronize()
        cold.append(a.elapsed_time(b) * 1e3)
    print("CONFIG1 cold single launches (us): " + " ".join(f"{x:.1f}" for x in cold))


if __name__ == "__main__":
    main()
