#!/bin/bash
# Round-2 job H: ncu launch list of the codec decode (bench shape) to see the per-stage attention kernel times
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 600 python scripts/profile_path.py --what codec > gpurun_out/r02h_plain.log 2>&1; echo "plain exit $?"; tail -5 gpurun_out/r02h_plain.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02h_codec_launches.csv python scripts/profile_path.py --what codec > gpurun_out/r02h_ncu.log 2>&1; echo "ncu exit $?"
python scripts/summarize_launches.py gpurun_out/r02h_codec_launches.csv 2>&1 | tail -40
