#!/bin/bash
# Round-2 job T: CTA-pair GEMM (gemm_pair.cu) - parity vs the one-SM kernel, then the codec / prefill timings with both
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 600 python -m pytest tests/test_gpu_gemm.py -m gpu -q -x --timeout 300 > gpurun_out/r02t_gemm.log 2>&1; echo "gemm tests exit $?"; tail -5 gpurun_out/r02t_gemm.log | cut -c1-300
timeout 900 python -m pytest tests/test_gpu_codec.py tests/test_gpu_codec_encode.py tests/test_gpu_baseline_shapes.py tests/test_gpu_lm.py -m gpu -q -x --timeout 600 > gpurun_out/r02t_tests.log 2>&1; echo "codec/lm tests exit $?"; tail -5 gpurun_out/r02t_tests.log | cut -c1-300
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r02t_bench.log 2> gpurun_out/r02t_bench.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r02t_bench.log').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value', 'ms_per_step', 'decode_ms_per_step', 'prefill_ms', 'codec_ms') if k in d}, d.get('e2e', {}).get('value'))
PY
