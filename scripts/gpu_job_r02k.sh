#!/bin/bash
# Round-2 job K: quick A/B of a codec change: attention tests + plain codec timing (3 repeats)
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 600 python -m pytest tests/test_gpu_codec.py -q -m gpu --timeout 600 -x -k "attention or golden" > gpurun_out/r02k_codec.log 2>&1; echo "codec exit $?"; tail -2 gpurun_out/r02k_codec.log
for i in 1 2 3; do timeout 600 python scripts/profile_path.py --what codec 2>&1 | tail -1; done
