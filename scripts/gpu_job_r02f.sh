#!/bin/bash
# Round-2 job F: tcgen05 codec attention tests (+ the sampling / lm / local tests after the oracle's float64 sums)
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 900 python -m pytest tests/test_gpu_codec.py -q -m gpu --timeout 600 -s -x > gpurun_out/r02f_codec.log 2>&1; echo "codec exit $?"
grep -E "passed|failed|^FAILED|\[attention|^E  " gpurun_out/r02f_codec.log | tail -30
timeout 1200 python -m pytest tests/test_gpu_sampling.py tests/test_gpu_lm.py tests/test_gpu_local.py tests/test_gpu_realtime.py -q -m gpu --timeout 900 -s > gpurun_out/r02f_tests.log 2>&1; echo "pytest exit $?"
grep -E "passed|failed|^FAILED|full-vocabulary|text sampling|^E  " gpurun_out/r02f_tests.log | tail -30
