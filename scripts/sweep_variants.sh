#!/bin/bash
# Tuning helper (GPU box): kernel-only bench of every lib/libnxs_variant_*.so x tile size; prints ms/step.
mkdir -p gpurun_out
for lib in nexus-supervisor_b200/lib/libnxs_variant_*.so; do
  tiles=${TILES:-2048}
  case "$lib" in *__*) tiles=$(basename $lib .so | sed 's/.*__//; s/_/ /g');; esac # libnxs_variant_w12__1536_3072.so
  for tile in $tiles; do
    out=$(NXS_VERBOSE=1 NXS_LIB=$PWD/$lib NXS_TILE=$tile timeout 120 python bench.py --no-e2e --no-cpu-baseline --steps 10 2> gpurun_out/sweep.err | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('%.4f ms  frac %.3f' % (d['ms_per_step'], d['roofline']['frac']))
except Exception as e: print('failed', e)")
    occ=$(grep -o "[0-9]* resident CTAs/SM x [0-9]* threads" gpurun_out/sweep.err | head -1)
    echo "SWEEP $(basename $lib) tile=$tile: $out  [$occ]"
  done
done
