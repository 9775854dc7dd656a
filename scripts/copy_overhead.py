#!/usr/bin/env python3
"""Fixed cost of one page-locked host<->device copy: back-to-back copies of several sizes on one stream,
alone and with the opposite direction busy on another stream.  Needs a GPU."""
import time
import torch

dev = torch.device("cuda:0")
big = 256 << 20
h_src = torch.empty(big, dtype=torch.uint8).pin_memory()
h_dst = torch.empty(big, dtype=torch.uint8).pin_memory()
d_a = torch.empty(big, dtype=torch.uint8, device=dev)
d_b = torch.empty(big, dtype=torch.uint8, device=dev)
s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()


def run(size, up, down):
    reps = max(4, min(256, big // size))
    best = 1e9
    for _ in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(reps):
            o = (i * size) % (big - size + 1)
            if up:
                with torch.cuda.stream(s_up):
                    d_a[o:o + size].copy_(h_src[o:o + size], non_blocking=True)
            if down:
                with torch.cuda.stream(s_dn):
                    h_dst[o:o + size].copy_(d_b[o:o + size], non_blocking=True)
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    return best / reps


for size in (8, 64 << 10, 256 << 10, 1 << 20, 2 << 20, 4 << 20, 8 << 20, 16 << 20, 64 << 20):
    u, d, b = run(size, True, False), run(size, False, True), run(size, True, True)
    print(f"{size:>10} B: h2d {u * 1e6:8.1f} us ({size / u / 1e9:5.1f} GB/s)  d2h {d * 1e6:8.1f} us ({size / d / 1e9:5.1f} GB/s)  "
          f"both {b * 1e6:8.1f} us ({size / b / 1e9:5.1f} GB/s each way)", flush=True)
