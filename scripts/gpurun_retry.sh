#!/usr/bin/env bash
# gpurun with retries while the pod answers "transient" (busy, nothing charged).  usage: gpurun_retry.sh [gpurun args] -- 'cmd'
for i in $(seq 1 30); do
  out=$(/usr/local/graft/bin/gpurun "$@" 2>&1)
  if echo "$out" | grep -q "status=transient"; then sleep 60; continue; fi
  echo "$out"; exit 0
done
echo "$out"; exit 3
