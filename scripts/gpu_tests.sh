#!/bin/bash
# Staged GPU validation: (1) everything on the SIMT cross-check GEMM, (2) the tcgen05 GEMM alone,
# (3) the whole suite on the tcgen05 GEMM.  Each stage has its own timeout so a hang cannot eat the box.
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
echo "== stage 1: SIMT-forced suite"; MT_FORCE_SIMT=1 timeout 1200 python -m pytest tests -q -m gpu -x --timeout 600 -s > gpurun_out/stage1_simt.log 2>&1; echo "exit $?"
tail -5 gpurun_out/stage1_simt.log
echo "== stage 2: tcgen05 GEMM"; timeout 600 python -m pytest tests/test_gpu_gemm.py -q -m gpu --timeout 300 > gpurun_out/stage2_gemm.log 2>&1; s2=$?; echo "exit $s2"
tail -15 gpurun_out/stage2_gemm.log
if [ $s2 -eq 0 ]; then
  echo "== stage 3: full suite on tcgen05"; timeout 1500 python -m pytest tests -q -m gpu --timeout 600 -s > gpurun_out/stage3_full.log 2>&1; echo "exit $?"
  tail -15 gpurun_out/stage3_full.log
fi
