#!/bin/bash
# usage: tools/gpurun_retry.sh <timeout-seconds> <command string> [--gpus N]
# retries while the pod answers "busy" (exit code 3); everything else is returned as is
T=$1; CMD=$2; shift 2
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" --timeout "$T" -- "$CMD"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 120
done
exit 3
