#!/bin/bash
# Round-2 job I: tcgen05 attention v2 (2 CTAs/SM, P over S, rotary epilogue in the in_proj GEMM): tests + launch list + bench
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 900 python -m pytest tests/test_gpu_codec.py tests/test_gpu_baseline_shapes.py tests/test_gpu_codec_encode.py tests/test_gpu_gemm.py -q -m gpu --timeout 600 -s -x > gpurun_out/r02i_codec.log 2>&1; echo "codec exit $?"
grep -E "passed|failed|^FAILED|\[attention|^E  |codec-full" gpurun_out/r02i_codec.log | tail -30
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02i_codec_launches.csv python scripts/profile_path.py --what codec > gpurun_out/r02i_ncu.log 2>&1; echo "ncu exit $?"
python scripts/summarize_launches.py gpurun_out/r02i_codec_launches.csv 2>&1 | grep -E "tc5|gemm_big|rope|layernorm|launches" | head
timeout 600 python scripts/profile_path.py --what codec 2>&1 | tail -2
