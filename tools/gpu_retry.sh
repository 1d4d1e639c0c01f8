#!/bin/bash
# tools/gpu_retry.sh <log> <gpurun args...>: retry while the pod answers "busy" (exit code 3, nothing charged)
log=$1; shift
for try in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  [ $rc -ne 3 ] && break
  sleep 90
done
echo "gpurun rc=$rc after $try tries" >> "$log"
exit $rc
