#!/bin/bash
# timing ablations of the tensor-core inverse (profiling build; results are garbage by construction)
mkdir -p gpurun_out
for a in ${ABL:-0 1 2 4 8 5 6 7}; do
  echo "ablate=$a"; SDCT_TC_ABLATE=$a timeout 120 python tools/prof_run.py --reps 20 inv_A_tensor inv_C_tensor 2>&1 | tail -2
done > gpurun_out/ablate_inv.txt 2>&1
awk '/ablate/{a=$0} /inv_A/{x=$2} /inv_C/{print a, "A", x, "C", $2}' gpurun_out/ablate_inv.txt
