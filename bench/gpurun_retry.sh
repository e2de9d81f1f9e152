#!/usr/bin/env bash
# Development helper: run a gpurun call, retrying while the pool answers "busy" (exit 3, nothing charged).
#   bench/gpurun_retry.sh <log file> [gpurun options] -- '<command>'
log=$1; shift
for attempt in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 90
done
exit 3
