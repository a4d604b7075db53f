#!/bin/bash
# usage: gpurun_retry.sh <timeout-seconds> <logfile> <command...>: retries while the pod answers busy / transient
T=$1; LOG=$2; shift 2
for attempt in $(seq 1 40); do
    /usr/local/graft/bin/gpurun --timeout "$T" -- "$@" > "$LOG" 2>&1
    if grep -q "status=transient\|status=busy\|retry in a few minutes\|rc=3" "$LOG" && ! grep -q "status=ok" "$LOG"; then sleep 45; continue; fi
    break
done
