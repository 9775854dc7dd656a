#!/usr/bin/env python3
"""Debug tool: runs one classification launch with the NXS_TRACE build of the library and prints, for a
few CTAs, how long each pipeline step of a tile takes (producer claim/issue, phase 1, aggregate,
look-back, phase 2).  Build first: make -C nexus-supervisor_b200/csrc trace.  Needs a GPU."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nexus_supervisor_b200 as nx  # noqa: E402
from nexus_supervisor_b200 import _abi  # noqa: E402

_abi.LIB_PATH = os.path.join(ROOT, "nexus-supervisor_b200", "lib", "libnxs_classify_trace.so")
import torch  # noqa: E402

EV = {1: "claim", 2: "issued", 3: "p1_start", 4: "p1_end", 5: "agg_start", 6: "agg_end", 7: "lb_start", 8: "lb_end",
      9: "p2_start", 10: "p2_end", 11: "lb_poll"}


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
    clf = nx.Classifier(device=0, max_records=n)
    lib = _abi.load()
    lib.nxs_debug_dump_trace.restype = C.c_longlong
    lib.nxs_debug_dump_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
    clf.set_stream(torch.cuda.current_stream().cuda_stream)
    cap = (n + 1023) & ~1023
    raw = torch.empty(cap * 13, dtype=torch.uint8, device="cuda")
    b = raw.data_ptr()
    ptr = dict(job_key=b, reason_id=b + cap * 8, kind=b + cap * 10, flags=b + cap * 11, cur_stage=b + cap * 12)
    clf.generate_device(nx.WORKLOAD_SKEWED, n, ptr["kind"], ptr["reason_id"], ptr["flags"], ptr["cur_stage"],
                        ptr["job_key"], cap)
    verdict = torch.empty(n, dtype=torch.int32, device="cuda")
    rows = torch.empty((n, 2), dtype=torch.int64, device="cuda")
    nj = torch.zeros(1, dtype=torch.int64, device="cuda")
    for _ in range(3):
        clf.classify_device(n, ptr["kind"], ptr["reason_id"], ptr["flags"], ptr["cur_stage"], ptr["job_key"],
                            verdict.data_ptr(), rows.data_ptr(), nj.data_ptr())
    torch.cuda.synchronize()
    buf = np.zeros(2 << 20, dtype=np.uint64)
    cnt = lib.nxs_debug_dump_trace(clf._ctx, buf.ctypes.data, 1 << 20)
    ev = buf[:2 * cnt].reshape(-1, 2)
    t0 = int(ev[:, 0].min())
    per = {}
    polls = {}
    for t, w in ev:
        cta, e, tile = int(w) >> 40, (int(w) >> 32) & 255, int(w) & 0xFFFFFFFF
        if e == 11:
            polls[(cta, tile)] = polls.get((cta, tile), 0) + 1
        else:
            per.setdefault((cta, tile), {})[EV[e]] = int(t) - t0
    print(f"{cnt} events, span {(int(ev[:, 0].max()) - t0) / 1e3:.1f} us")
    for cta in sorted({k[0] for k in per}):
        tiles = sorted(k[1] for k in per if k[0] == cta)
        print(f"--- CTA {cta}: {len(tiles)} tiles; times in us relative to kernel start; durations in ns")
        print(" tile    claim  | load  wait  P1   gap  AGG  gap  LB(polls)  gap  P2 | claim->P2end")
        agg = {k: [] for k in ("load", "w1", "p1", "g1", "agg", "g2", "lb", "g3", "p2", "tot", "polls")}
        for i, tile in enumerate(tiles):
            d = per[(cta, tile)]
            need = ("claim", "issued", "p1_start", "p1_end", "agg_start", "agg_end", "lb_start", "lb_end", "p2_start", "p2_end")
            if not all(k in d for k in need):
                continue
            vals = dict(load=d["issued"] - d["claim"], w1=d["p1_start"] - d["issued"], p1=d["p1_end"] - d["p1_start"],
                        g1=d["agg_start"] - d["p1_end"], agg=d["agg_end"] - d["agg_start"], g2=d["lb_start"] - d["agg_end"],
                        lb=d["lb_end"] - d["lb_start"], g3=d["p2_start"] - d["lb_end"], p2=d["p2_end"] - d["p2_start"],
                        tot=d["p2_end"] - d["claim"], polls=polls.get((cta, tile), 0))
            for k, v in vals.items():
                agg[k].append(v)
            if i < 12 or i % 40 == 0:
                print(f"{tile:6d} {d['claim'] / 1e3:8.1f} | {vals['load']:5d} {vals['w1']:5d} {vals['p1']:5d} {vals['g1']:5d} "
                      f"{vals['agg']:4d} {vals['g2']:5d} {vals['lb']:5d}({vals['polls']:2d}) {vals['g3']:5d} {vals['p2']:4d} | {vals['tot']:6d}")
        print(" median: " + "  ".join(f"{k}={int(np.median(v))}" for k, v in agg.items() if v))
        print(" mean:   " + "  ".join(f"{k}={int(np.mean(v))}" for k, v in agg.items() if v))
        if len(tiles) > 1:
            claims = np.array([per[(cta, t)]["claim"] for t in tiles if "claim" in per[(cta, t)]])
            print(f" tile period (claim to claim): median {int(np.median(np.diff(np.sort(claims))))} ns")


if __name__ == "__main__":
    main()
