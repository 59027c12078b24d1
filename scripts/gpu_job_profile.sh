#!/bin/bash
# Round measurement job (run through gpurun): tests, bench, ncu launch lists and full captures -> gpurun_out/<tag>_*
tag=${1:-r01g}
o=gpurun_out
mkdir -p $o
export PYTHONDONTWRITEBYTECODE=1
python -m pytest tests -m gpu -q 2>&1 | tail -5 > $o/${tag}_tests.log
python bench.py > $o/${tag}_bench.log 2> $o/${tag}_bench.err
python scripts/bench_sampling.py > $o/${tag}_sampling.jsonl 2> $o/${tag}_sampling.err
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
# codec: the fp16 GEMM (a stage-3 linear1) and the fp16 attention (stage 3), one launch each
ncu --set full --clock-control none --import-source on -k regex:codec_attention_f16_kernel -s 60 -c 1 -f -o $o/${tag}_codec_attn \
    python scripts/profile_path.py --what codec > $o/${tag}_ncufull_cattn.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_big_kernel -s 262 -c 2 -f -o $o/${tag}_codec_gemm \
    python scripts/profile_path.py --what codec > $o/${tag}_ncufull_cgemm.log 2>&1
for r in gemm_dec attn codec_attn codec_gemm; do
  python scripts/ncu_full_summary.py $o/${tag}_$r.ncu-rep "$tag ncu --set full: $r" > $o/${tag}_${r}_ncu_full.md 2>/dev/null
done
tail -3 $o/${tag}_tests.log; cat $o/${tag}_bench.log
