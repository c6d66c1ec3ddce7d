#!/bin/bash
# usage: tools/gpurun_retry.sh LOGFILE TIMEOUT 'command' [extra gpurun flags, e.g. --gpus 2]
# retries while the pod answers "busy" (exit 3)
log=$1; shift; to=$1; shift; cmd=$1; shift
for k in $(seq 1 20); do
  /usr/local/graft/bin/gpurun --timeout $to "$@" -- "$cmd" > $log 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 45
done
exit 3
