#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "sse or dct8_bitwise or inv255 or config1 or reference" 2>&1 | tail -6
timeout 300 python tools/prof_run.py --reps 10 fwd_C_exact 2>&1 | tail -1
for spec in "fwd_B_exact forward_kernel" "fwd_B_tensor forward_tc_kernel" "inv_B_tensor inverse_tc_kernel" "rt_B_tensor roundtrip_tc_kernel" "fwd_A_tensor forward_tc_kernel"; do
  set -- $spec
  timeout 300 bash tools/gpu_ncu.sh $1 $2 --reps 4 $1 > /dev/null 2>&1
  ls -la gpurun_out/$1_hot.txt
done
