#!/bin/bash
# usage: scripts/gpurun_retry.sh <out-file> <gpurun args...>   — retries while the pod answers busy (exit code 3)
OUT=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > "$OUT" 2>&1; rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" "$OUT"; then exit $rc; fi
  sleep 120
done
exit 3
