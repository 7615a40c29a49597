#!/bin/bash
# timing ablations of the tensor-core kernel (results are garbage by construction; profiling only)
mkdir -p gpurun_out
for a in ${ABL:-0 1 2 4 8 16 18 19 5 20 21 23 31}; do
  echo "ablate=$a"; SDCT_TC_ABLATE=$a timeout 120 python tools/prof_run.py --reps 20 fwd_A_tensor fwd_B_tensor 2>&1 | tail -2
done > gpurun_out/ablate.txt 2>&1
awk '/ablate/{a=$0} /fwd_A/{x=$2} /fwd_B/{print a, "A", x, "B", $2}' gpurun_out/ablate.txt
