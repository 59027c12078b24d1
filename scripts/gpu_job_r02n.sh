#!/bin/bash
# Round-2 job N: early-read decode attention (product build): LM-side GPU tests twice (races show up as flakiness), bench
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
for i in 1 2; do
timeout 1500 python -m pytest tests/test_gpu_lm.py tests/test_gpu_local.py tests/test_gpu_realtime.py tests/test_gpu_full_size.py tests/test_gpu_baseline_shapes.py tests/test_gpu_stream_synth.py -q -m gpu --timeout 900 > gpurun_out/r02n_tests$i.log 2>&1; echo "tests exit $?"; tail -2 gpurun_out/r02n_tests$i.log
done
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r02n_bench.json 2> gpurun_out/r02n_bench.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r02n_bench.json').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value', 'ms_per_step', 'decode_ms_per_step', 'prefill_ms', 'codec_ms') if k in d}, d.get('e2e', {}).get('value'), d['roofline']['frac'], d['roofline_step']['frac'])
PY
timeout 600 python scripts/bench_configs.py --what rt,c4 2>/dev/null | cut -c1-330
