#!/usr/bin/env python3
"""Where the samples and the issued instructions of an ncu capture fall, by source line and by stage.

  python scripts/ncu_hotspots.py REPORT.ncu-rep LIB.so KERNEL_SUBSTR [--stages spec.json] [--top N]

`ncu --page source --print-source sass` gives per-SASS-instruction sample counts without line numbers;
`nvdisasm -g` of the same library gives the line table.  Both list the kernel's instructions in the
same order, so they are joined by position.  --stages maps source-line ranges to names as in
scripts/sass_stage_table.py.
"""
import argparse
import collections
import csv
import io
import json
import os
import re
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from sass_stage_table import disassemble  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("lib")
    ap.add_argument("kernel")
    ap.add_argument("--stages")
    ap.add_argument("--top", type=int, default=25)
    args = ap.parse_args()
    stages = json.load(open(args.stages)) if args.stages else {}

    out = subprocess.run(["ncu", "-i", args.report, "--page", "source", "--csv", "--print-source", "sass"],
                         capture_output=True, text=True).stdout
    lines = out.splitlines()
    start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
    rd = list(csv.DictReader(io.StringIO("\n".join(lines[start:]))))
    sass = [(r["Source"].strip(), int(r["# Samples"] or 0), int(r["Instructions Executed"] or 0), r) for r in rd]

    text = disassemble(args.lib)
    in_fn, cur, dis = False, ("?", 0), []
    for line in text.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", line)
        if m:
            in_fn = args.kernel in m.group(1) and "$" not in m.group(1)
            continue
        if not in_fn:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', line)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m:
            dis.append((cur, m.group(2).strip()))
    n = min(len(dis), len(sass))
    if len(dis) != len(sass):
        print(f"# note: {len(sass)} instructions in the report, {len(dis)} in the kernel's main body of the library "
              "(out-of-line functions follow the main body in the report)")

    def stage_of(key):
        for name, ranges in stages.items():
            for f, lo, hi in ranges:
                if key[0] == f and lo <= key[1] <= hi:
                    return name
        return f"{key[0]}:{key[1]}"

    samples, issued = collections.Counter(), collections.Counter()
    by_line_s, by_line_i = collections.Counter(), collections.Counter()
    stall = collections.defaultdict(collections.Counter)
    for i, (src, smp, ex, r) in enumerate(sass):
        key = dis[i][0] if i < n else ("out-of-line", 0)
        st = stage_of(key)
        samples[st] += smp
        issued[st] += ex
        by_line_s[key] += smp
        by_line_i[key] += ex
        for k, v in r.items():
            if k.startswith("stall_") and "Not Issued" not in k and v:
                stall[st][k[6:]] += int(v)
    ts, ti = sum(samples.values()), sum(issued.values())
    print(f"# {ts} samples, {ti} warp-instructions issued")
    print("# stage: % of samples, % of issued instructions, top stall reasons")
    for st, c in samples.most_common(args.top):
        top = ", ".join(f"{k} {v * 100 // max(1, c)}%" for k, v in stall[st].most_common(4))
        print(f"{c * 100.0 / ts:5.1f}%  {issued[st] * 100.0 / ti:5.1f}%  {st}   [{top}]")
    print("# hottest source lines")
    for key, c in by_line_s.most_common(args.top):
        print(f"{c * 100.0 / ts:5.1f}%  {by_line_i[key] * 100.0 / ti:5.1f}%  {key[0]}:{key[1]}")


if __name__ == "__main__":
    main()
