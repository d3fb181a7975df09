#!/bin/bash
# usage: scripts/gpurun_retry.sh LOGFILE [gpurun args...]; retries while the pod answers "busy" (rc 3), up to ~40 min
LOG=$1; shift
for i in $(seq 1 14); do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 170
done
exit 3
