#!/bin/bash
# retry_gpurun.sh <log> <timeout-seconds> [--gpus N] -- '<command>' : gpurun, retried while it answers "no box free" (rc 3)
log=$1; to=$2; shift 2
extra=()
while [ "$1" != "--" ]; do extra+=("$1"); shift; done
shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$to" "${extra[@]}" -- "$@" > "$log" 2>&1
  rc=$?
  echo "attempt $i rc=$rc" >> "$log.attempts"
  [ $rc -ne 3 ] && exit $rc
  sleep 90
done
exit 3
