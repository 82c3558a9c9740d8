#!/usr/bin/env bash
# usage: gpurun_retry.sh <timeout> <command string>   — retries while the pod answers busy (exit 3)
T="$1"; shift
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --timeout "$T" -- "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 60
done
exit 3
