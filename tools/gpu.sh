#!/bin/bash
# Rebuild the CUDA library, then run a command on a B200 through gpurun.
#   tools/gpu.sh <timeout-seconds> '<command>'
set -e
cd "$(dirname "$0")/.."
python -m imaginaire_b200.build > /dev/null
python -c "from oracle import build_oracle; build_oracle.build()"
exec /usr/local/graft/bin/gpurun --timeout "$1" -- "$2"
