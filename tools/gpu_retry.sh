#!/bin/bash
# Developer helper: retry a gpurun call while the pod answers "busy" (exit code 3).  usage: tools/gpu_retry.sh <timeout> '<command>'
T=$1; shift
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --timeout "$T" -- "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 150
done
exit 3
