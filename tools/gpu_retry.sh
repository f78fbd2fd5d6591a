#!/bin/bash
# gpurun with retries while the pod answers "busy / draining" (exit code 3, nothing charged)
# usage: tools/gpu_retry.sh <timeout-seconds> <command...>
T=$1; shift
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --timeout "$T" -- "$@" > /tmp/gpu_retry.out 2>&1
  rc=$?
  if grep -q "nothing was charged" /tmp/gpu_retry.out; then sleep 90; continue; fi
  break
done
tail -12 /tmp/gpu_retry.out
exit $rc
