#!/bin/bash
# usage: tools/gpurun_retry.sh <logfile> <gpurun args...>   -- repeats a gpurun call while the pod answers "transient"
# (no box / slot free, nothing charged), every 2 minutes, at most 30 times
LOG=$1; shift
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
  if ! grep -q "status=transient" "$LOG"; then exit 0; fi
  sleep 120
done
