#!/bin/bash
# usage: gpurun_retry.sh <timeout> [--gpus N] -- '<command>'   (retries while the pod answers "busy", exit code 3)
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 120
done
exit 3
