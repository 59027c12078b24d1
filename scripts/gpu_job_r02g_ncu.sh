#!/bin/bash
# ncu --set full of the token-major GEMMs after the round-2 epilogue / pair-kernel work
tag=${1:-r02g}
o=gpurun_out
mkdir -p $o
export PYTHONDONTWRITEBYTECODE=1
# prefill: gate|up and down GEMMs of a middle layer now run on gemm_pair_kernel<1> (4 GEMM launches per layer)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_pair_kernel -s 42 -c 2 -f -o $o/${tag}_prefill_gemm \
    python scripts/profile_path.py --what lm --steps 2 > $o/${tag}_ncufull_pgemm.log 2>&1; echo "ncu prefill gemm exit $?"
# codec: one-SM kernel (in_proj / out_proj / linear1: 3 per layer) of a stage-3 layer, and the pair kernel (linear2)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_big_kernel -s 186 -c 3 -f -o $o/${tag}_codec_gemm \
    python scripts/profile_path.py --what codec > $o/${tag}_ncufull_cgemm.log 2>&1; echo "ncu codec gemm exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_pair_kernel -s 62 -c 1 -f -o $o/${tag}_codec_gemm_pair \
    python scripts/profile_path.py --what codec > $o/${tag}_ncufull_cgemmp.log 2>&1; echo "ncu codec gemm pair exit $?"
for r in prefill_gemm codec_gemm codec_gemm_pair; do
  python scripts/ncu_full_summary.py $o/${tag}_$r.ncu-rep "$tag ncu --set full: $r" > $o/${tag}_${r}_ncu_full.md 2>/dev/null
done
cat $o/${tag}_*_ncu_full.md | cut -c1-300
