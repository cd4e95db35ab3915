#!/bin/bash
# usage: tools/gpurun_retry.sh <logfile> <timeout> [--gpus N] <command>   - retries while the pod answers busy (nothing charged)
log=$1; shift; to=$1; shift
extra=()
if [ "$1" = "--gpus" ]; then extra=(--gpus "$2"); shift 2; fi
for attempt in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$to" "${extra[@]}" -- "$@" > "$log" 2>&1
  rc=$?
  if grep -q "status=transient\|retry in a few minutes\|status=busy" "$log" || [ $rc -eq 3 ]; then sleep 150; continue; fi
  exit $rc
done
exit 3
