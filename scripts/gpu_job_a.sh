#!/bin/bash
# GPU job: parity suite, bench line, standalone sampler / RVQ rooflines -> gpurun_out/<tag>_*
tag=${1:-r01c}
o=gpurun_out
mkdir -p $o
export PYTHONDONTWRITEBYTECODE=1
( time timeout 1200 python -m pytest tests -m gpu -q --timeout 600 ) > $o/${tag}_tests.log 2>&1
tail -5 $o/${tag}_tests.log
( time timeout 600 python bench.py ) > $o/${tag}_bench.log 2> $o/${tag}_bench.err
cat $o/${tag}_bench.log; tail -3 $o/${tag}_bench.err
timeout 300 python scripts/bench_sampling.py > $o/${tag}_sampling.jsonl 2> $o/${tag}_sampling.err
cat $o/${tag}_sampling.jsonl; tail -3 $o/${tag}_sampling.err
