#!/bin/bash
# usage: bash scripts/gpu_quick.sh <pytest args...>   (logs to gpurun_out/quick.log)
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 1500 python -m pytest "$@" -q -m gpu --timeout 900 -s > gpurun_out/quick.log 2>&1
echo "pytest exit $?"
tail -40 gpurun_out/quick.log
