#!/usr/bin/env python3
"""Launch latency of nxs_classify_device for small and medium batches (device-resident, CUDA events),
per built-in tile size (NXS_TILE).  Two figures per size: back-to-back relaunches over the same buffers
(which stay in the 126 MB L2 below a few million records) and COLD launches — a 512 MB buffer is
rewritten before every launch, so inputs come from HBM.  Needs a GPU."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def one(tile):
    os.environ["NXS_TILE"] = tile
    import torch
    import nexus_supervisor_b200 as nx
    sizes = [int(x) for x in os.environ.get("NXS_SIZES", "1000,10000,100000,1000000,4000000,10000000").split(",")]
    cap = max(sizes) + 4096
    clf = nx.Classifier(device=0, max_records=cap)
    stream = torch.cuda.current_stream()
    clf.set_stream(stream.cuda_stream)
    raw = torch.empty(cap * 13, dtype=torch.uint8, device="cuda")
    b = raw.data_ptr()
    ptr = dict(job_key=b, reason_id=b + cap * 8, kind=b + cap * 10, flags=b + cap * 11, cur_stage=b + cap * 12)
    clf.generate_device(nx.WORKLOAD_SKEWED, cap, ptr["kind"], ptr["reason_id"], ptr["flags"], ptr["cur_stage"],
                        ptr["job_key"], cap)
    verdict = torch.empty(cap, dtype=torch.int32, device="cuda")
    rows = torch.empty((cap, 2), dtype=torch.int64, device="cuda")
    nj = torch.zeros(1, dtype=torch.int64, device="cuda")
    out = []
    flush = torch.zeros(128 << 20, dtype=torch.int32, device="cuda")
    for n in sizes:
        def step():
            clf.classify_device(n, ptr["kind"], ptr["reason_id"], ptr["flags"], ptr["cur_stage"], ptr["job_key"],
                                verdict.data_ptr(), rows.data_ptr(), nj.data_ptr())
        for _ in range(5):
            step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 50
        e0.record(stream)
        for _ in range(reps):
            step()
        e1.record(stream)
        torch.cuda.synchronize()
        warm = e0.elapsed_time(e1) / reps * 1e3
        cold = []
        for _ in range(12):
            flush.add_(1)  # 512 MB read + written: evicts the batch from L2
            e0.record(stream)
            step()
            e1.record(stream)
            torch.cuda.synchronize()
            cold.append(e0.elapsed_time(e1) * 1e3)
        out.append((warm, float(sorted(cold)[len(cold) // 2])))
    print(f"TILE {tile:>5}: " + "  ".join(f"n={n:>9,}: {w:7.1f} / {c:7.1f} us" for n, (w, c) in zip(sizes, out)) +
          "   (relaunch / cold median)")


if __name__ == "__main__":
    if len(sys.argv) > 1:
        one(sys.argv[1])
    else:
        for t in ("2048", "4096", "8192"):
            subprocess.run([sys.executable, __file__, t], check=False)
