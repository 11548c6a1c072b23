#!/bin/bash
# usage: gpurun_retry.sh <logfile> <timeout> <gpus> <command>   (retries while the pod answers busy / transient)
log=$1; shift; to=$1; shift; gp=$1; shift
for i in $(seq 1 40); do
  if [ "$gp" = "1" ]; then gpurun --timeout $to -- "$@" > $log 2>&1; else gpurun --gpus $gp --timeout $to -- "$@" > $log 2>&1; fi
  if grep -q "status=transient\|status=busy\|nothing was charged" $log; then echo "retry $i $(date)" >> $log.retries; sleep 120; else break; fi
done
echo finished >> $log.retries
