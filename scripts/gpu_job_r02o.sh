#!/bin/bash
# Round-2 job O (EXPERIMENTS build): delayed early ring fill of the decode GEMMs (x 256 ns)
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
for f in 0 4 8 16 0 8; do
  echo "== MT_DEC_DELAY=$f"
  MT_DEC_DELAY=$f timeout 600 python scripts/trace_decode.py --steps 60 2>&1 | grep -E "decode ms/step|^\| o \||^\| gate|^\| down|^\| qkv" | cut -c1-120 | head -5
done
