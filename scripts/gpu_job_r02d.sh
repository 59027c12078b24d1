#!/bin/bash
# Round-2 job D (2 GPUs): NCCL gather test, N=2 bench line (NCCL gather of padded waveforms inside the e2e step).
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
echo "== nccl gather test + changed tests"; timeout 900 python -m pytest tests/test_gpu_sharding_nccl.py tests/test_gpu_stream_synth.py tests/test_gpu_sampling.py tests/test_gpu_lm.py -q -m gpu --timeout 600 -s > gpurun_out/r02d_tests.log 2>&1; echo "exit $?"
grep -E "passed|failed|^E  |natural|Error" gpurun_out/r02d_tests.log | tail -15
echo "== bench N=2"; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29631 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r02d_bench_n2.log 2> gpurun_out/r02d_bench_n2.err; echo "exit $?"
tail -c 3000 gpurun_out/r02d_bench_n2.log; tail -3 gpurun_out/r02d_bench_n2.err
echo "== smoke"; timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
