#!/bin/bash
# usage: tools/gpurun_retry.sh TIMEOUT 'command'   -- retries while the pod answers "busy / draining"
T=$1; shift
for i in $(seq 1 20); do
  out=$(/usr/local/graft/bin/gpurun --timeout "$T" -- "$@" 2>&1)
  echo "$out" | tail -40
  if echo "$out" | grep -q "status=transient\|status=busy\|exit code 3"; then sleep 240; continue; fi
  break
done
