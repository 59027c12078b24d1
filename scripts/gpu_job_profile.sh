#!/bin/bash
# Round measurement job (run through gpurun): tests, bench, ncu launch lists and full captures -> gpurun_out/<tag>_*
tag=${1:-r01b}
o=gpurun_out
mkdir -p $o
python -m pytest tests -m gpu -q 2>&1 | tail -5 > $o/${tag}_tests.log
python bench.py > $o/${tag}_bench.log 2> $o/${tag}_bench.err
python scripts/profile_path.py --steps 6 > $o/${tag}_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $o/${tag}_launches.csv \
    python scripts/profile_path.py --steps 6 > $o/${tag}_ncu.log 2>&1
python bench.py --steps 1 --warmup 1 --no-cpu > $o/${tag}_bench_short.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 110000 -c 1500 --csv --log-file $o/${tag}_bench_window.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu > $o/${tag}_ncu_bench.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_dec_kernel -s 150 -c 4 -f -o $o/${tag}_gemm_dec \
    python scripts/profile_path.py --what lm --steps 4 > $o/${tag}_ncufull_gemm.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:attn_decode_kernel -s 40 -c 1 -f -o $o/${tag}_attn \
    python scripts/profile_path.py --what lm --steps 4 > $o/${tag}_ncufull_attn.log 2>&1
tail -3 $o/${tag}_tests.log; cat $o/${tag}_bench.log
