#!/bin/bash
# usage: tools/gpurun_retry.sh <timeout-seconds> '<command>' [--gpus N]  -- gpurun, retried while the pod answers "busy" (exit 3)
T=$1; CMD=$2; shift 2
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun "$@" --timeout "$T" -- "$CMD" > /tmp/gpurun_last.txt 2>&1
  rc=$?
  cat /tmp/gpurun_last.txt
  if [ $rc -ne 3 ] && ! grep -q 'status=transient' /tmp/gpurun_last.txt; then exit $rc; fi
  sleep 90
done
exit 3
