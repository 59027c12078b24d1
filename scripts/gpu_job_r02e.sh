#!/bin/bash
# Round-2 job E: re-run the sampling / lm / local tests after the oracle's large-candidate float64 sums, then the decode trace.
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 1200 python -m pytest tests/test_gpu_sampling.py tests/test_gpu_lm.py tests/test_gpu_local.py tests/test_gpu_realtime.py -q -m gpu --timeout 900 -s > gpurun_out/r02e_tests.log 2>&1; echo "pytest exit $?"
grep -E "passed|failed|^FAILED|full-vocabulary|text sampling" gpurun_out/r02e_tests.log | tail -30
timeout 600 python scripts/trace_decode.py --steps 40 > gpurun_out/r02e_trace.log 2>&1; echo "trace exit $?"
tail -45 gpurun_out/r02e_trace.log
