#!/bin/bash
# Here (no GPU): gpurun with retries while the pod answers "busy" (exit 3, nothing charged).
# Usage: scripts/grun.sh <log> <timeout-seconds> '<command>'
log=$1; to=$2; shift 2
for try in $(seq 1 40); do
  gpurun --timeout $to -- "$@" > $log 2>&1
  rc=$?
  if grep -q "status=transient" $log; then sleep 45; continue; fi
  exit $rc
done
exit 3
