#!/bin/bash
# Dev helper (container side): run a gpurun call, retrying while the pod answers "busy" (exit 3).
# Usage: bash scripts/gpurun_retry.sh <timeout-seconds> '<command>' [--gpus N]
T=$1; CMD=$2; shift 2
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" --timeout "$T" -- "$CMD"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 90
done
exit 3
