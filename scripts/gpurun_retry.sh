#!/bin/bash
# usage: scripts/gpurun_retry.sh <log> <timeout> [--gpus N] -- <command>: retries while the pod answers busy (exit 3 / transient)
LOG=$1; TO=$2; shift 2
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $TO "$@" > $LOG 2>&1
  if grep -q "status=transient" $LOG; then sleep 100; continue; fi
  break
done
