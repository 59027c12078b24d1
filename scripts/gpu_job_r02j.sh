#!/bin/bash
# Round-2 job J: ncu --set full of one stage-3 launch of the tcgen05 codec attention
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:codec_attention_tc5 -s 60 -c 1 -f -o gpurun_out/r02j_tc5 python scripts/profile_path.py --what codec > gpurun_out/r02j_ncu.log 2>&1; echo "ncu exit $?"
ls -la gpurun_out/r02j_tc5.ncu-rep
