#!/bin/bash
# Run on the B200 box via gpurun: kernel-only bench, ncu launch list + one full capture of the classification
# kernel, worker idle/busy totals (stats build) and the per-tile pipeline trace (trace build).  Logs under gpurun_out/.
set -u
mkdir -p gpurun_out
BARGS="--steps 3 --warmup 3 --no-e2e --no-cpu-baseline --records 100000000"
echo "== bench (kernel only)"; timeout 300 python bench.py --no-e2e --no-cpu-baseline > gpurun_out/bench_k.json 2> gpurun_out/bench_k.err; echo "rc=$?"
python - <<PY
import json
d = json.load(open("gpurun_out/bench_k.json"))
print("%.4f ms/step  %.1f GB/s  frac %.3f  value %.3e" % (d["ms_per_step"], d["roofline"]["achieved"], d["roofline"]["frac"], d["value"]))
PY
if [ "${SKIP_NCU:-0}" != "1" ]; then
echo "== ncu launch list"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches.csv python bench.py $BARGS > gpurun_out/ncu_launches.log 2>&1; echo "rc=$?"
echo "== ncu full"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:classify_kernel -s 3 -c 1 -f -o gpurun_out/prof python bench.py $BARGS > gpurun_out/ncu_full.log 2>&1; echo "rc=$?"
fi
if [ -f nexus-supervisor_b200/lib/libnxs_variant_stats.so ]; then
echo "== worker idle/busy"; NXS_TRACE_LIB=$PWD/nexus-supervisor_b200/lib/libnxs_variant_stats.so timeout 300 python scripts/trace_pipeline.py > gpurun_out/worker_idle.txt 2>&1; echo "rc=$?"; head -4 gpurun_out/worker_idle.txt | cut -c1-400
fi
if [ -f nexus-supervisor_b200/lib/libnxs_classify_trace.so ]; then
echo "== pipeline trace"; timeout 300 python scripts/trace_pipeline.py > gpurun_out/pipeline_trace.txt 2>&1; echo "rc=$?"; grep -E "median|period" gpurun_out/pipeline_trace.txt | head -12
fi
