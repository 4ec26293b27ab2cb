#!/bin/bash
# Retry a gpurun call while the pod answers "busy" (exit 3: nothing charged).  Usage:
#   tools/gpurun_retry.sh <log> [gpurun args...] -- '<command>'
log=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc after $i tries" >> "$log"; exit $rc; fi
  sleep 120
done
echo "gpurun: still busy after 40 tries" >> "$log"; exit 3
