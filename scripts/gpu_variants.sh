#!/bin/bash
# Kernel-only bench of "lib:tile" pairs given in PAIRS (B200 box).
mkdir -p gpurun_out
for pt in $PAIRS; do
  lib=${pt%%:*}; t=${pt#*:}
  NXS_VERBOSE=1 NXS_LIB=$PWD/nexus-supervisor_b200/lib/$lib NXS_TILE=$t timeout 300 python bench.py --no-cpu-baseline --no-e2e --steps ${STEPS:-10} > gpurun_out/bt.json 2> gpurun_out/bt.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bt.json"))
    print("VAR $lib tile $t: %.4f ms/step  frac %.3f  value %.3e" % (d["ms_per_step"], d["roofline"]["frac"], d["value"]))
except Exception as e:
    print("VAR $lib tile $t: no result", e); print(open("gpurun_out/bt.err").read()[-300:])
PY
done
