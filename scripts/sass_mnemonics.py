#!/usr/bin/env python3
"""Mnemonic counts of the classification kernels in a built library (cuobjdump -sass):
  python scripts/sass_mnemonics.py [nexus-supervisor_b200/lib/libnxs_classify.so]"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "nexus-supervisor_b200", "lib", "libnxs_classify.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
cur, counts = None, collections.defaultdict(collections.Counter)
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and cur:
        counts[cur][m.group(1)] += 1
WANT = ["UBLKCP", "UBLKPF", "SYNCS", "LDGSTS", "ARRIVES", "VOTE", "VOTEU", "SHFL", "SHF", "FLO", "POPC", "R2P", "LDS", "STS",
        "STG", "LDG", "ATOMG", "NANOSLEEP", "LOP3", "IMAD", "BAR", "ELECT", "CCTL", "MEMBAR", "FENCE", "HMMA", "UTCMMA"]
for fn, c in sorted(counts.items()):
    if "classify_kernel" not in fn:
        continue
    print(fn)
    for w in WANT:
        if c[w]:
            print("  %-10s %d" % (w, c[w]))
    print("  total      %d" % sum(c.values()))
