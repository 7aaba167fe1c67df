#!/bin/sh
# usage: tools/gpurun_retry.sh <timeout_s> '<command>'   -- retries while the pod answers busy (nothing charged)
T=$1; shift
for i in $(seq 1 40); do
    /usr/local/graft/bin/gpurun --timeout "$T" -- "$@"
    st=$(python -c "import json;print(json.load(open('/root/repo/gpurun_out/.last_call.json'))['status'])" 2>/dev/null)
    [ "$st" = "transient" ] || exit 0
    echo "[retry $i] pod busy; sleeping 90 s"
    sleep 90
done
