#!/bin/bash
# Round-2 job P (8 GPUs): request-sharded records of BASELINE configs 4 and 5, and the N = 8 bench line
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
: > gpurun_out/r02p_c45.jsonl
for N in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29700 + N)) scripts/bench_configs.py --what c4 >> gpurun_out/r02p_c45.jsonl 2> gpurun_out/r02p_c4_n$N.err; echo "c4 N=$N exit $?"
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29720 scripts/bench_configs.py --what c5 >> gpurun_out/r02p_c45.jsonl 2> gpurun_out/r02p_c5_n8.err; echo "c5 N=8 exit $?"
cut -c1-360 gpurun_out/r02p_c45.jsonl
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29730 bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/r02p_bench_n8.json 2> gpurun_out/r02p_bench_n8.err; echo "bench N=8 exit $?"
python - <<'PY'
import json
try:
    d = json.loads(open('gpurun_out/r02p_bench_n8.json').read().strip().splitlines()[-1])
    print({k: d[k] for k in ('value', 'n_gpus', 'ms_per_step', 'decode_ms_per_step') if k in d}, d.get('e2e', {}))
except Exception as e:
    print('parse failed', e)
PY
