#!/bin/bash
# usage: tools/grun_retry.sh <gpus> <timeout_s> <logfile> '<command>'   -- gpurun with retries while the pod is busy
G=$1; T=$2; LOG=$3; shift 3
for i in 1 2 3 4 5 6 7 8 9 10; do
  /usr/local/graft/bin/gpurun --gpus $G --timeout $T -- "$@" > $LOG 2>&1
  if grep -q "status=transient\|status=busy\|rc=3\|exit code 3\|no box" $LOG; then sleep 120; continue; fi
  break
done
