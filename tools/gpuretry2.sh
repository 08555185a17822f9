#!/bin/bash
# usage: tools/gpuretry2.sh <logfile> <ngpus> <timeout-seconds> <command>   -- multi-GPU gpurun, retried while the pod is busy
log=$1; shift; n=$1; shift; to=$1; shift
for i in $(seq 1 60); do
  /usr/local/graft/bin/gpurun --gpus $n --timeout $to -- "$@" > $log 2>&1
  if grep -q "status=transient\|status=busy\|rc=3" $log && ! grep -q "status=ok" $log; then sleep 150; continue; fi
  break
done
