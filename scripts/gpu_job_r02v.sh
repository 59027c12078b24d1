#!/bin/bash
# Round-2 job V: token-major GEMM epilogues (warp-transposed stores, fast GELU), CTA-pair kernel - parity, A/B, bench
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 600 python -m pytest tests/test_gpu_gemm.py -m gpu -q -x --timeout 300 > gpurun_out/r02v_gemm.log 2>&1; echo "gemm tests exit $?"; tail -5 gpurun_out/r02v_gemm.log | cut -c1-300
timeout 600 python scripts/ab_gemm_pair.py > gpurun_out/r02v_ab.jsonl 2> gpurun_out/r02v_ab.err; echo "ab exit $?"
timeout 1200 python -m pytest tests/test_gpu_codec.py tests/test_gpu_codec_encode.py tests/test_gpu_codec_stream.py tests/test_gpu_baseline_shapes.py tests/test_gpu_lm.py tests/test_gpu_full_size.py -m gpu -q -x --timeout 600 -s 2>&1 | grep -E "SNR|passed|failed|Error|error" | cut -c1-250 > gpurun_out/r02v_tests.log; echo "codec/lm tests exit $?"; tail -30 gpurun_out/r02v_tests.log
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r02v_bench.log 2> gpurun_out/r02v_bench.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r02v_bench.log').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value', 'ms_per_step', 'decode_ms_per_step', 'prefill_ms', 'codec_ms') if k in d}, d.get('e2e', {}).get('value'))
PY
