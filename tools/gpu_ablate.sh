#!/bin/bash
# timing ablations of the tensor-core kernel (results are garbage by construction; profiling only)
mkdir -p gpurun_out
for a in 0 1 2 4 8 16 3 5 20 21 23 31; do
  echo "ablate=$a"; SDCT_TC_ABLATE=$a timeout 120 python tools/prof_run.py --reps 20 fwd_A_tensor fwd_B_tensor 2>&1 | tail -2
done > gpurun_out/ablate.txt 2>&1
cat gpurun_out/ablate.txt
