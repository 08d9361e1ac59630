#!/bin/bash
# gpurun with retries while the pod answers busy / transient (rc 3 or status=transient); usage: gpurun_retry.sh <log> <timeout> <command...>
log=$1; shift; to=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $to -- "$@" > $log 2>&1; rc=$?
  if grep -q "status=transient\|no box\|busy" $log || [ $rc -eq 3 ]; then sleep 120; continue; fi
  break
done
exit $rc
