#!/bin/bash
# usage: gpurun_retry.sh <logfile> <gpurun args...>   — retries while the pod answers "busy / draining" (nothing charged)
LOG=$1; shift
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
  if grep -q "status=transient\|status=busy\|rc=3" "$LOG" && ! grep -q "charged=[1-9]" "$LOG"; then sleep 90; continue; fi
  break
done
echo finished >> "$LOG"
