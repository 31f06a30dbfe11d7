#!/bin/bash
# gpurun with retries while the pod answers busy / transient (exit 3).  usage: tools/gpurun_retry.sh <log> <gpurun args...>
LOG=$1; shift
for k in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "transient" "$LOG"; then echo "DONE rc=$rc" >> "$LOG"; exit $rc; fi
  sleep 90
done
echo "DONE gave up" >> "$LOG"
