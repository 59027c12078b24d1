#!/bin/bash
# Round-2 measurement job (run through gpurun): tests, bench, ncu launch lists and --set full captures -> gpurun_out/<tag>_*
tag=${1:-r02}
o=gpurun_out
mkdir -p $o
export PYTHONDONTWRITEBYTECODE=1
timeout 2400 python -m pytest tests -m gpu -q --timeout 900 > $o/${tag}_tests.log 2>&1; echo "tests exit $?"; tail -3 $o/${tag}_tests.log
timeout 1500 python bench.py > $o/${tag}_bench.log 2> $o/${tag}_bench.err; echo "bench exit $?"
timeout 600 python scripts/bench_sampling.py > $o/${tag}_sampling.jsonl 2> $o/${tag}_sampling.err; echo "sampling exit $?"
timeout 600 python scripts/profile_path.py --steps 6 > $o/${tag}_plain.log 2>&1 && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $o/${tag}_launches.csv \
    python scripts/profile_path.py --steps 6 > $o/${tag}_ncu.log 2>&1; echo "launch list exit $?"
timeout 900 python bench.py --steps 1 --warmup 3 --no-cpu > $o/${tag}_bench_short.log 2>&1 && \
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -s 110000 -c 1500 --csv --log-file $o/${tag}_bench_window.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu > $o/${tag}_ncu_bench.log 2>&1; echo "bench window exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_dec_kernel -s 150 -c 4 -f -o $o/${tag}_gemm_dec \
    python scripts/profile_path.py --what lm --steps 4 > $o/${tag}_ncufull_gemm.log 2>&1; echo "ncu gemm_dec exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_decode_kernel -s 40 -c 1 -f -o $o/${tag}_attn_decode \
    python scripts/profile_path.py --what lm --steps 4 > $o/${tag}_ncufull_attn.log 2>&1; echo "ncu attn_decode exit $?"
# prefill: the bf16 gate|up GEMM of a middle layer (gemm_big_kernel<1>: launches 0.. are prefill GEMMs, 4 per layer)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_big_kernel -s 42 -c 2 -f -o $o/${tag}_prefill_gemm \
    python scripts/profile_path.py --what lm --steps 2 > $o/${tag}_ncufull_pgemm.log 2>&1; echo "ncu prefill gemm exit $?"
# codec: the tcgen05 attention (a stage-3 launch) and the fp16 GEMM (stage-3 linear1 / linear2)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:codec_attention_tc5 -s 60 -c 1 -f -o $o/${tag}_codec_attn_tc5 \
    python scripts/profile_path.py --what codec > $o/${tag}_ncufull_cattn.log 2>&1; echo "ncu tc5 exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_big_kernel -s 262 -c 2 -f -o $o/${tag}_codec_gemm \
    python scripts/profile_path.py --what codec > $o/${tag}_ncufull_cgemm.log 2>&1; echo "ncu codec gemm exit $?"
for r in gemm_dec attn_decode prefill_gemm codec_attn_tc5 codec_gemm; do
  python scripts/ncu_full_summary.py $o/${tag}_$r.ncu-rep "$tag ncu --set full: $r" > $o/${tag}_${r}_ncu_full.md 2>/dev/null
done
python scripts/summarize_launches.py $o/${tag}_launches.csv > $o/${tag}_launch_summary.md 2>&1
cat $o/${tag}_bench.log | cut -c1-600
