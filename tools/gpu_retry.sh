#!/bin/bash
# Like tools/gpu.sh, but retries while the pod answers "busy / draining" (exit code 3).
#   tools/gpu_retry.sh <timeout-seconds> '<command>'
cd "$(dirname "$0")/.."
python -m imaginaire_b200.build > /dev/null || exit 1
python -c "from oracle import build_oracle; build_oracle.build()" || exit 1
for attempt in $(seq 1 20); do
    /usr/local/graft/bin/gpurun --timeout "$1" -- "$2"
    rc=$?
    if [ $rc -ne 3 ]; then exit $rc; fi
    sleep 150
done
exit 3
