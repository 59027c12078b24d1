#!/bin/bash
# Round-2 job L: full GPU suite on the current build, Realtime / C3 / encode / C4 measurements
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 2400 python -m pytest tests -q -m gpu --timeout 900 > gpurun_out/r02l_tests.log 2>&1; echo "pytest exit $?"
tail -6 gpurun_out/r02l_tests.log
timeout 900 python scripts/bench_configs.py --what rt > gpurun_out/r02l_rt.jsonl 2> gpurun_out/r02l_rt.err; echo "rt exit $?"; cat gpurun_out/r02l_rt.jsonl | cut -c1-400; tail -3 gpurun_out/r02l_rt.err
timeout 900 python scripts/bench_configs.py --what c3 > gpurun_out/r02l_c3.jsonl 2> gpurun_out/r02l_c3.err; echo "c3 exit $?"; tail -3 gpurun_out/r02l_c3.jsonl | cut -c1-300
timeout 900 python scripts/bench_configs.py --what enc > gpurun_out/r02l_enc.jsonl 2> gpurun_out/r02l_enc.err; echo "enc exit $?"; tail -2 gpurun_out/r02l_enc.jsonl | cut -c1-300
timeout 900 python scripts/bench_configs.py --what c4 > gpurun_out/r02l_c4.jsonl 2> gpurun_out/r02l_c4.err; echo "c4 exit $?"; cat gpurun_out/r02l_c4.jsonl | cut -c1-400
