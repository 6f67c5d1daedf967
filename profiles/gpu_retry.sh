#!/bin/bash
# Usage: profiles/gpu_retry.sh LOGFILE [gpurun args] -- 'command'   retries while the pod answers "busy" (exit 3)
LOG=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
  rc=$?
  [ $rc -ne 3 ] && exit $rc
  sleep 120
done
exit 3
