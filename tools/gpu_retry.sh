#!/bin/bash
# tools/gpu_retry.sh <log> <timeout-seconds> <command...>: run one gpurun call, retrying while the pod answers "busy" (rc 3).
LOG=$1; shift; TMO=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $TMO -- "$@" > $LOG 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" $LOG; then exit $rc; fi
  sleep 90
done
exit 3
