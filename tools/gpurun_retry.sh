#!/bin/bash
# tools/gpurun_retry.sh TIMEOUT 'command' [--gpus N] -- retries a gpurun call while the pod answers "busy" (exit 3).
T=$1; CMD=$2; shift 2
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun --timeout "$T" "$@" -- "$CMD"; rc=$?
  [ $rc -ne 3 ] && exit $rc
  sleep 150
done
exit 3
