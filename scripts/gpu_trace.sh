#!/bin/bash
# Worker idle/busy totals and per-tile pipeline trace for the tile sizes in TILES (B200 box).
mkdir -p gpurun_out
for t in ${TILES:-4096}; do
  echo "== tile $t: worker idle/busy"
  NXS_TILE=$t NXS_TRACE_LIB=$PWD/nexus-supervisor_b200/lib/libnxs_variant_stats.so timeout 300 python scripts/trace_pipeline.py > gpurun_out/worker_idle_$t.txt 2>&1; head -3 gpurun_out/worker_idle_$t.txt | cut -c1-330
  echo "== tile $t: pipeline trace"
  NXS_TILE=$t timeout 300 python scripts/trace_pipeline.py > gpurun_out/pipeline_trace_$t.txt 2>&1; grep -E "median|period" gpurun_out/pipeline_trace_$t.txt | head -4
done
