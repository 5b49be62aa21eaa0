#!/bin/bash
# usage: tools/gpurun_retry.sh <logfile> <gpurun args...>   -- retries while the pod has no free GPU slot (nothing is charged for those)
log=$1; shift
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  if grep -q "status=transient\|retry in a few minutes\|busy" "$log" && ! grep -q "charged=[1-9]" "$log"; then sleep 150; continue; fi
  break
done
echo "[retry wrapper] finished after $i attempt(s)" >> "$log"
