#!/bin/bash
# usage: gpu_ncu.sh <outname> <kernel-regex> <prof_run names...>
out=$1; shift; rx=$1; shift
mkdir -p gpurun_out
python tools/prof_run.py "$@" > gpurun_out/${out}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$rx -s 2 -c 2 -f -o gpurun_out/$out python tools/prof_run.py "$@" > gpurun_out/${out}_ncu.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/${out}_plain.log; tail -3 gpurun_out/${out}_ncu.log
# the summary csv travels back (gpurun_out/ is capped at 64 MiB); the report itself only with KEEP_REP=1
python tools/ncu_summary.py gpurun_out/$out.ncu-rep gpurun_out/${out}_ncu.csv 2>/dev/null
# per-instruction warp-stall samples (SASS view) -> the 40 hottest instructions with their dominant stall reasons
ncu -i gpurun_out/$out.ncu-rep --page source --csv > gpurun_out/${out}_source.csv 2>/dev/null &&
  python tools/ncu_hot.py gpurun_out/${out}_source.csv 40 > gpurun_out/${out}_hot.txt 2>/dev/null
rm -f gpurun_out/${out}_source.csv
[ -n "$KEEP_REP" ] || rm -f gpurun_out/$out.ncu-rep
