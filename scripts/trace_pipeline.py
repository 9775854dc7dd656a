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

_abi.LIB_PATH = os.environ.get("NXS_TRACE_LIB") or os.path.join(ROOT, "nexus-supervisor_b200", "lib", "libnxs_classify_trace.so")
import torch  # noqa: E402

EV = {1: "claim", 2: "issued", 3: "count_start", 4: "count_end", 5: "lb_start", 6: "lb_end", 7: "p1_start",
      8: "p1_end", 9: "sum_start", 10: "sum_end", 11: "lb_poll", 12: "slb_end"}


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
    steps = {}
    stats = {}
    for t, w in ev:
        cta, e, tile = int(w) >> 40, (int(w) >> 32) & 255, int(w) & 0xFFFFFFFF
        if e == 11:
            polls[(cta, tile)] = polls.get((cta, tile), 0) + 1
        elif e in (14, 15, 18):  # NXS_TRACE_STATS build: per-worker-warp idle / busy totals (ns / 16)
            stats.setdefault(cta, {}).setdefault(tile >> 28, {})[{14: "idle", 15: "busy", 18: "idle_base"}[e]] = (tile & 0xFFFFFFF) * 16
        elif e == 13:  # NXS_TRACE_STEPS build: end of 32-record step j of warp 0
            steps.setdefault((cta, tile & 0xFFFFFF), {})[tile >> 24] = int(t) - t0
        else:
            per.setdefault((cta, tile), {})[EV[e]] = int(t) - t0
    print(f"{cnt} events, span {(int(ev[:, 0].max()) - t0) / 1e3:.1f} us")
    for cta, ws in sorted(stats.items()):
        print(f"--- CTA {cta}: worker warps, us waiting (of which for the first row index) / us classifying: " +
              "  ".join(f"w{w}: {v.get('idle', 0) / 1e3:.0f}({v.get('idle_base', 0) / 1e3:.0f})/{v.get('busy', 0) / 1e3:.0f}"
                        for w, v in sorted(ws.items())))
    for cta in sorted({k[0] for k in per}):
        tiles = sorted(k[1] for k in per if k[0] == cta)
        print(f"--- CTA {cta}: {len(tiles)} tiles; claim in us from kernel start; durations in ns")
        print(" tile    claim  | issue  land+wait count  gap   LB(polls)  wait   P1    gap  SUM | claim->stage free")
        cols = ("issue", "land", "count", "g1", "lb", "w2", "p1", "g2", "sum", "hold", "polls")
        agg = {k: [] for k in cols}
        for i, tile in enumerate(tiles):
            d = per[(cta, tile)]
            need = ("claim", "issued", "count_start", "count_end", "lb_start", "lb_end", "p1_start", "p1_end",
                    "sum_start", "sum_end")
            if not all(k in d for k in need):
                continue
            vals = dict(issue=d["issued"] - d["claim"], land=d["count_start"] - d["issued"],
                        count=d["count_end"] - d["count_start"], g1=d["lb_start"] - d["count_end"],
                        lb=d["lb_end"] - d["lb_start"], w2=d["p1_start"] - d["lb_end"], p1=d["p1_end"] - d["p1_start"],
                        g2=d["sum_start"] - d["p1_end"], sum=d["sum_end"] - d["sum_start"],
                        hold=d["sum_start"] - d["claim"], polls=polls.get((cta, tile), 0))
            for k, v in vals.items():
                agg[k].append(v)
            if i < 12 or i % 40 == 0:
                print(f"{tile:6d} {d['claim'] / 1e3:8.1f} | " + " ".join(f"{vals[k]:6d}" for k in cols))
        print(" median: " + "  ".join(f"{k}={int(np.median(v))}" for k, v in agg.items() if v))
        print(" mean:   " + "  ".join(f"{k}={int(np.mean(v))}" for k, v in agg.items() if v))
        chain = np.array([per[(cta, t)]["lb_end"] - per[(cta, t)]["claim"] for t in tiles
                          if "lb_end" in per[(cta, t)] and "claim" in per[(cta, t)]])
        mk = np.array([per[(cta, t)]["count_end"] - per[(cta, t)]["claim"] for t in tiles
                       if "count_end" in per[(cta, t)] and "claim" in per[(cta, t)]])
        if len(chain):
            q = lambda a: " / ".join(str(int(x)) for x in np.percentile(a, [50, 90, 99, 100]))  # noqa: E731
            print(f" claim -> count published (ns, median / p90 / p99 / max): {q(mk)};  claim -> first row index known: {q(chain)}")
        st = [(per[k]["p1_start"], per[k].get("p1_end"), steps[k]) for k in steps if k[0] == cta and "p1_start" in per.get(k, {})]
        if st:
            nj = max(len(x[2]) for x in st)
            first = np.median([x[2][0] - x[0] for x in st if 0 in x[2]])
            later = [np.median([x[2][j] - x[2][j - 1] for x in st if j in x[2] and j - 1 in x[2]]) for j in range(1, nj)]
            tail = np.median([x[1] - x[2][nj - 1] for x in st if x[1] and nj - 1 in x[2]])
            print(f" steps of warp 0 (ns): start->step0 {int(first)}  step j-1->j {[int(v) for v in later]}  last step->end {int(tail)}")
        seq = sorted((per[(cta, t)]["p1_start"], per[(cta, t)].get("p1_end"), per[(cta, t)].get("issued"), t) for t in tiles
                     if "p1_start" in per[(cta, t)])
        gaps = [b[0] - a[1] for a, b in zip(seq, seq[1:]) if a[1]]
        land = [x[0] - x[2] for x in seq if x[2]]
        if gaps:
            print(f" workers: end of a tile's passes -> start of the next record pass (ns, median / p90 / max): "
                  f"{int(np.median(gaps))} / {int(np.percentile(gaps, 90))} / {int(max(gaps))};  bulk copies issued -> record pass "
                  f"starts: {int(np.median(land))} / {int(np.percentile(land, 90))} / {int(max(land))}")
        if len(tiles) > 1:
            claims = np.array([per[(cta, t)]["claim"] for t in tiles if "claim" in per[(cta, t)]])
            print(f" tile period (claim to claim): median {int(np.median(np.diff(np.sort(claims))))} ns")


if __name__ == "__main__":
    main()
