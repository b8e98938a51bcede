#!/bin/bash
# usage: scripts/gpu_retry.sh <logfile> <gpurun args...>   -- retries while the pod answers "busy" (rc 3)
log=$1; shift
for i in $(seq 1 40); do
  gpurun "$@" > "$log" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc" >> "$log"; exit $rc; fi
  sleep 90
done
echo "gpurun gave up" >> "$log"
