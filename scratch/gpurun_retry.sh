#!/bin/bash
# usage: scratch/gpurun_retry.sh <outfile> <gpurun args...>   - retries while the pod answers "busy / draining" (nothing charged)
out=$1; shift
for attempt in $(seq 1 40); do
  gpurun "$@" > "$out" 2>&1
  rc=$?
  if grep -q "status=transient\|retry in a few minutes" "$out" || [ $rc -eq 3 ]; then sleep 90; continue; fi
  break
done
echo "done rc=$rc attempt=$attempt" >> "$out"
