#!/bin/bash
# gpurun with retries on "no slot right now" (exit 3): tools/gpurun_retry.sh <log> <timeout> [--gpus N] -- <command>
log=$1; shift; to=$1; shift
for attempt in 1 2 3 4 5 6 7 8; do
  /usr/local/graft/bin/gpurun --timeout $to "$@" > $log 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 150
done
exit 3
