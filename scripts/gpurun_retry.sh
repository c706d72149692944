#!/bin/bash
# usage: scripts/gpurun_retry.sh <log> <timeout_s> [--gpus N] -- <command>     retries while the pod answers "busy" (exit 3)
LOG=$1; TMO=$2; shift 2
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun --timeout $TMO "$@" > $LOG 2>&1
  rc=$?
  [ $rc -ne 3 ] && exit $rc
  sleep 100
done
exit 3
