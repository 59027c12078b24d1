#!/bin/bash
# Round-2 job M (EXPERIMENTS build): decode attention with position-dependent loads hoisted above the dependency wait
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 900 python -m pytest tests/test_gpu_lm.py tests/test_gpu_local.py tests/test_gpu_realtime.py tests/test_gpu_full_size.py -q -m gpu --timeout 600 -x > gpurun_out/r02m_tests.log 2>&1; echo "tests exit $?"; tail -2 gpurun_out/r02m_tests.log
for f in 5 1 9 5 1 9; do
  echo "== MT_ATTN_EXP=$f (bit0: no post-wait L2 prefetch, bit2: no hoisting, bit3: pre-wait K/V L2 prefetch)"
  MT_ATTN_EXP=$f timeout 600 python scripts/trace_decode.py --steps 60 2>&1 | grep -E "decode ms/step|^\| o \||^\| qkv" | head -4
done
