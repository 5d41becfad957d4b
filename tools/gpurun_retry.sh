#!/bin/bash
# Retry a gpurun call while the pod answers "busy" (exit 3: nothing charged) or another call of this repo is still in flight (exit 2).
# Usage: tools/gpurun_retry.sh [gpurun args] -- 'cmd'
for i in $(seq 1 40); do
  out=$(/usr/local/graft/bin/gpurun "$@" 2>&1)
  rc=$?
  echo "$out"
  if [ $rc -eq 2 ] && echo "$out" | grep -q "already running"; then sleep 60; continue; fi
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 150
done
exit 3
