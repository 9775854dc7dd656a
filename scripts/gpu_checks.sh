#!/bin/bash
# Run on the B200 box via gpurun: smoke, GPU parity tests, bench, then the ncu launch list and one
# full capture of the classification kernel.  Everything is logged under gpurun_out/.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
echo "== smoke" ; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1 ; echo "smoke rc=$?" ; tail -3 gpurun_out/smoke.log
echo "== pytest gpu" ; timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1 ; echo "pytest rc=$?" ; tail -15 gpurun_out/pytest_gpu.log
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
