#!/bin/bash
# usage: tools/gpuretry.sh <logfile> <timeout-seconds> <command>   -- gpurun, retried while the pod answers "transient / busy"
log=$1; shift; to=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $to -- "$@" > $log 2>&1
  if grep -q "status=transient" $log; then sleep 120; continue; fi
  break
done
