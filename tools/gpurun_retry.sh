#!/bin/bash
# gpurun with retries while the pool answers "no box free" (exit 3, nothing charged).
#   tools/gpurun_retry.sh <timeout-seconds> [--gpus N] -- '<command>'
T=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$T" "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 90
done
exit 3
