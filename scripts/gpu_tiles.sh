#!/bin/bash
# Kernel-only bench per tile size (and per NXS_LIB variant given in LIBS), on the B200 box.
mkdir -p gpurun_out
for lib in ${LIBS:-nexus-supervisor_b200/lib/libnxs_classify.so}; do
for t in ${TILES:-2048 4096 8192}; do
  NXS_LIB=$PWD/$lib NXS_TILE=$t timeout 300 python bench.py --no-cpu-baseline --no-e2e --steps ${STEPS:-20} > gpurun_out/bt.json 2> gpurun_out/bt.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bt.json"))
    print("TILE $t $(basename $lib): %.4f ms/step  %.1f GB/s  frac %.3f  value %.3e" % (d["ms_per_step"], d["roofline"]["achieved"], d["roofline"]["frac"], d["value"]))
except Exception as e:
    print("TILE $t $(basename $lib): no result", e)
PY
done
done
