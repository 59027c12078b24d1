#!/bin/bash
# Round-2 job G: sampling tests again, baseline-shape goldens (68-layer codec through the tcgen05 attention), bench line
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 1200 python -m pytest tests/test_gpu_sampling.py tests/test_gpu_lm.py tests/test_gpu_local.py tests/test_gpu_baseline_shapes.py tests/test_gpu_codec_stream.py tests/test_gpu_codec_encode.py -q -m gpu --timeout 900 -s > gpurun_out/r02g_tests.log 2>&1; echo "pytest exit $?"
grep -E "passed|failed|^FAILED|full-vocabulary|^E  |dB" gpurun_out/r02g_tests.log | tail -40
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/r02g_bench.json 2> gpurun_out/r02g_bench.err; echo "bench exit $?"
python - <<'PY'
import json
try:
    d = json.loads(open('gpurun_out/r02g_bench.json').read().strip().splitlines()[-1])
    print({k: d[k] for k in ('value', 'ms_per_step', 'decode_ms_per_step', 'prefill_ms', 'codec_ms') if k in d}, d.get('e2e', {}).get('value'))
except Exception as e:
    print('bench parse failed', e); print(open('gpurun_out/r02g_bench.err').read()[-2000:])
PY
