#!/bin/bash
# usage: tools/gpurun_retry.sh <timeout-seconds> '<command>'   - retries while the pod answers busy / transient
t=$1; shift
for i in $(seq 1 20); do
  out=$(/usr/local/graft/bin/gpurun --timeout "$t" -- "$@" 2>&1); rc=$?
  echo "$out" | tail -4
  if echo "$out" | grep -q "status=transient\|status=busy\|retry in a few minutes"; then sleep 90; continue; fi
  exit $rc
done
exit 3
