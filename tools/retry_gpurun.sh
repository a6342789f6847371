#!/bin/bash
# tools/retry_gpurun.sh LOG TRIES GPURUN-ARGS... : retry a gpurun call while the pod answers busy (exit code 3 / backing off)
log=$1; tries=$2; shift; shift
for i in $(seq 1 $tries); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  if grep -q "status=ok\|status=failed\|status=timeout" "$log"; then exit 0; fi
  if [ $rc -eq 2 ] && ! grep -q "backing off" "$log"; then exit 2; fi
  sleep 240
done
exit 3
