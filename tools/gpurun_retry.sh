#!/bin/bash
# usage: tools/gpurun_retry.sh <log> <gpus> <timeout_s> <command...>   -- retries while the pod answers busy / transient (rc 3)
log=$1; gpus=$2; to=$3; shift 3
: > "$log.attempts"
for i in $(seq 1 40); do
  if [ "$gpus" = 1 ]; then /usr/local/graft/bin/gpurun --timeout "$to" -- "$@" > "$log" 2>&1; else /usr/local/graft/bin/gpurun --gpus "$gpus" --timeout "$to" -- "$@" > "$log" 2>&1; fi
  rc=$?
  echo "attempt $i rc=$rc $(date +%T)" >> "$log.attempts"
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 45
done
exit 3
