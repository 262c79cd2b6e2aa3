#!/bin/bash
# developer helper: gpurun with retries while the pod answers "transient" (nothing charged); usage: gpu_retry.sh LOG TIMEOUT 'command'
LOG=$1; TMO=$2; shift 2
for k in $(seq 1 20); do
  /usr/local/graft/bin/gpurun --timeout "$TMO" -- "$@" > "$LOG" 2>&1
  if ! grep -q "status=transient" "$LOG"; then exit 0; fi
  sleep 45
done
exit 3
