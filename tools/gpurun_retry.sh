#!/bin/bash
# gpurun with retries while the pod answers "transient" (busy): tools/gpurun_retry.sh LOGFILE TIMEOUT [--gpus N] -- 'command'
LOG="$1"; shift; TMO="$1"; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$TMO" "$@" > "$LOG" 2>&1
  if grep -q "status=transient\|status=busy" "$LOG" || grep -q "retry in a few minutes" "$LOG"; then sleep 45; continue; fi
  break
done
tail -5 "$LOG"
