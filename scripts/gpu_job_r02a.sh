#!/bin/bash
# Round-2 job A: full GPU suite (with printed parity numbers), bench line, reference-on-GPU record.
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
nproc > gpurun_out/nproc.txt; free -g >> gpurun_out/nproc.txt
echo "== tests"; timeout 1500 python -m pytest tests -q -m gpu --timeout 900 -s > gpurun_out/r02a_tests.log 2>&1; echo "exit $?"
grep -E "^\[|passed|failed|Error|error" gpurun_out/r02a_tests.log | tail -60
echo "== bench"; timeout 900 python bench.py > gpurun_out/r02a_bench.log 2> gpurun_out/r02a_bench.err; echo "exit $?"
tail -c 6000 gpurun_out/r02a_bench.log; tail -5 gpurun_out/r02a_bench.err
echo "== reference on GPU"; timeout 900 python scripts/reference_gpu.py > gpurun_out/r02a_reference_gpu.json 2> gpurun_out/r02a_reference_gpu.err; echo "exit $?"
cat gpurun_out/r02a_reference_gpu.json; tail -5 gpurun_out/r02a_reference_gpu.err
