#!/bin/bash
# usage: scripts/gpurun_retry.sh <log> <timeout> <command...>   -- retries while the pod has no free GPU slot (exit 3)
log=$1; shift; to=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $to -- "$@" > $log 2>&1
  rc=$?
  echo "exit $rc (attempt $i)" >> $log
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 150
done
