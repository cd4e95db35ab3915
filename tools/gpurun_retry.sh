#!/bin/bash
# usage: tools/gpurun_retry.sh <logfile> <timeout> <command...>   - retries while the pod answers "transient"/busy (nothing charged)
log=$1; shift; to=$1; shift
for attempt in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$to" -- "$@" > "$log" 2>&1
  rc=$?
  if grep -q "status=transient\|retry in a few minutes" "$log" || [ $rc -eq 3 ]; then sleep 150; continue; fi
  exit $rc
done
exit 3
