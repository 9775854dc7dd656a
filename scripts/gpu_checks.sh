#!/bin/bash
# Run on the B200 box via gpurun: smoke, GPU parity tests, bench, then the ncu launch list and one
# full capture of the classification kernel.  Everything is logged under gpurun_out/.
# Env: SKIP_TESTS=1, SKIP_NCU=1, TILES="2048 4096 8192" (extra kernel-only bench lines per tile variant)
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
echo "== smoke" ; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1 ; rc=$? ; echo "smoke rc=$rc" ; tail -3 gpurun_out/smoke.log
if [ $rc -ne 0 ]; then echo "smoke failed: stopping"; exit 1; fi
if [ "${SKIP_TESTS:-0}" != "1" ]; then
  echo "== pytest gpu" ; timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1 ; echo "pytest rc=$?" ; tail -15 gpurun_out/pytest_gpu.log
fi
for t in ${TILES:-}; do
  echo "== bench tile $t" ; NXS_TILE=$t timeout 600 python bench.py --no-cpu-baseline --no-e2e > gpurun_out/bench_tile$t.json 2> gpurun_out/bench_tile$t.err ; echo "rc=$?"
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_tile$t.json"))
    print("tile $t: %.3f ms/step  %.1f GB/s  frac %.3f  value %.3e" % (d["ms_per_step"], d["roofline"]["achieved"], d["roofline"]["frac"], d["value"]))
except Exception as e:
    print("tile $t: no result", e)
PY
done
echo "== bench" ; timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err ; echo "bench rc=$?" ; cat gpurun_out/bench.json ; tail -5 gpurun_out/bench.err
if [ "${SKIP_NCU:-0}" != "1" ]; then
  BARGS="--steps 3 --warmup 3 --no-e2e --no-cpu-baseline --records 100000000"
  echo "== ncu launch list"
  timeout 600 python bench.py $BARGS > gpurun_out/ncu_plain.log 2>&1 && \
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches.csv python bench.py $BARGS > gpurun_out/ncu_launches.log 2>&1
  echo "launch list rc=$?"
  echo "== ncu full"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:classify_kernel -s 3 -c 1 -f -o gpurun_out/prof python bench.py $BARGS > gpurun_out/ncu_full.log 2>&1
  echo "ncu full rc=$?"
fi
