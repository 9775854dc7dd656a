#!/bin/bash
# gpurun wrapper: retries only when no box was available (status=transient); prints the log tail.
# usage: scripts/gpu.sh <logfile> <command...>
log=$1; shift
for i in 1 2 3 4 5 6 7 8; do
  timeout 3300 /usr/local/graft/bin/gpurun --timeout 2400 -- "$@" > "$log" 2>&1
  if ! grep -q "status=transient" "$log"; then break; fi
  sleep 75
done
grep -E "rc=|passed|failed|status=" "$log" | head -12
