#!/bin/bash
# usage: scripts/gpurun_retry.sh <timeout-seconds> <command...>; retries while the pod answers "transient" (nothing charged)
T=$1; shift
for attempt in $(seq 1 60); do
  /usr/local/graft/bin/gpurun --timeout "$T" -- "$@" > /tmp/gpurun_last.log 2>&1
  if grep -q "status=transient" /tmp/gpurun_last.log; then sleep 45; continue; fi
  break
done
cat /tmp/gpurun_last.log
