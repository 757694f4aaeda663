#!/bin/bash
# usage: tools/grun.sh <logfile> <timeout_s> <command...>   -- retries while the pod is busy (nothing is charged then)
log=$1; shift; to=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $to -- "$@" > "$log" 2>&1
  if grep -q "status=transient\|status=busy\|rc=3\|retry in a few minutes" "$log" && ! grep -q "status=ok" "$log"; then sleep 45; continue; fi
  break
done
