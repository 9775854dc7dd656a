#!/usr/bin/env python3
"""Per-stage instruction table of a SASS address range (e.g. one unrolled step of the classification loop).

  python scripts/sass_stage_table.py LIB.so KERNEL_SUBSTR START END [--stages spec.json] [--lines]

Extracts the cubin with cuobjdump, disassembles it with `nvdisasm -g` (the library is built with
-lineinfo), and attributes every instruction in [START, END) of the kernel whose mangled name contains
KERNEL_SUBSTR to the innermost source (file, line) the line table gives.  `--stages` maps
source-line ranges to stage names: {"stage": [["nxs_kernel.cuh", lo, hi], ...], ...}; lines no range
covers are listed under their own file:line.  Static counts only: an instruction in a branch that few
lanes take still costs the warp one issue slot whenever any lane takes it.
"""
import argparse
import collections
import json
import os
import re
import subprocess
import tempfile


def disassemble(lib):
    tmp = tempfile.mkdtemp(prefix="nxs_sass_")
    subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, stdout=subprocess.DEVNULL)
    out = []
    for f in sorted(os.listdir(tmp)):
        if f.endswith(".cubin"):
            r = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], capture_output=True, text=True)
            out.append(r.stdout)
    return "\n".join(out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("lib")
    ap.add_argument("kernel")
    ap.add_argument("start", type=lambda s: int(s, 16))
    ap.add_argument("end", type=lambda s: int(s, 16))
    ap.add_argument("--stages")
    ap.add_argument("--lines", action="store_true", help="also print the per-line counts")
    args = ap.parse_args()
    stages = json.load(open(args.stages)) if args.stages else {}

    text = disassemble(args.lib)
    in_fn, cur = False, ("?", 0)
    per_line = collections.Counter()
    per_op = collections.Counter()
    listing = []
    for line in text.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", line)
        if m:
            in_fn = args.kernel in m.group(1)
            continue
        if not in_fn:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', line)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if not m:
            continue
        addr = int(m.group(1), 16)
        if not (args.start <= addr < args.end):
            continue
        ins = m.group(2).strip()
        op = re.sub(r"^@!?U?P\d+\s+", "", ins).split()[0].split(".")[0]
        per_line[cur] += 1
        per_op[op] += 1
        listing.append((addr, cur, ins))

    def stage_of(key):
        for name, ranges in stages.items():
            for f, lo, hi in ranges:
                if key[0] == f and lo <= key[1] <= hi:
                    return name
        return f"{key[0]}:{key[1]}"

    per_stage = collections.Counter()
    for key, c in per_line.items():
        per_stage[stage_of(key)] += c
    total = sum(per_line.values())
    print(f"# {args.kernel} [{args.start:#x}, {args.end:#x}): {total} instructions")
    for name, c in sorted(per_stage.items(), key=lambda kv: -kv[1]):
        print(f"{c:5d}  {name}")
    print("# by opcode: " + ", ".join(f"{op} {c}" for op, c in per_op.most_common()))
    if args.lines:
        for (f, ln), c in sorted(per_line.items()):
            print(f"    {f}:{ln}  {c}")


if __name__ == "__main__":
    main()
