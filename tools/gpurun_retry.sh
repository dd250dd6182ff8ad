#!/bin/bash
# gpurun with retries while the pod answers "busy / transient" (exit code 3, nothing charged).
# Usage: tools/gpurun_retry.sh <log> <gpurun args...>
LOG=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
  rc=$?
  if grep -q "status=transient\|retry in a few minutes" "$LOG" || [ $rc -eq 3 ]; then sleep 90; continue; fi
  break
done
echo "gpurun rc=$rc (attempt $i)" >> "$LOG"
