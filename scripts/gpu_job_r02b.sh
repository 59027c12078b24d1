#!/bin/bash
# Round-2 job B: the allocation-free entry point's test alone (fresh process), then the whole GPU suite (it runs last).
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
echo "== alone"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu --timeout 300 > gpurun_out/r02b_noalloc_alone.log 2>&1; echo "exit $?"
tail -5 gpurun_out/r02b_noalloc_alone.log
echo "== full suite"; timeout 1500 python -m pytest tests -x -q -m gpu --timeout 900 -s > gpurun_out/r02b_tests.log 2>&1; echo "exit $?"
grep -E "^\[(lm32|lm16|codec-full)|teacher-forced argmax|passed|failed|Error" gpurun_out/r02b_tests.log | tail -30
echo "== smoke"; timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
