#!/bin/bash
# ncu evidence for the bench command (B200_PROFILING.md recipe): launch list with device times, then a
# full-set capture of the dominant kernel.  Numbers printed under ncu are never bench values.
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
CMD="python bench.py --steps 1 --warmup 3 --no-cpu"
$CMD > gpurun_out/prof_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/prof_plain.log; exit 1; }
tail -c 600 gpurun_out/prof_plain.log; echo
SKIP=${SKIP:-209200}
ncu --metrics gpu__time_duration.sum --clock-control none -s $SKIP -c 1400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "ncu launch list exit $?"; wc -l gpurun_out/launches.csv
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 2000 -c 4 -o gpurun_out/prof_gemm $CMD > gpurun_out/ncu2.log 2>&1
echo "ncu full exit $?"; ls -la gpurun_out/*.ncu-rep
