#!/bin/bash
# usage: tools/gpu_retry.sh <logfile> <gpurun args...>   retries while the pod answers "busy" (exit code 3), 2 min apart
log=$1; shift
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc (attempt $i)" >> "$log"; exit $rc; fi
  sleep 120
done
echo "gave up" >> "$log"
