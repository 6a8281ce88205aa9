#!/bin/bash
# usage: tools/gpurun_retry.sh <logfile> <gpurun args...>   -- retries while the pod answers "busy" (exit 3)
LOG=$1; shift
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun "$@" > $LOG 2>&1
  rc=$?
  echo "exit $rc (try $i)" >> $LOG
  if [ $rc -ne 3 ]; then break; fi
  sleep 90
done
