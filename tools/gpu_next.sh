#!/bin/bash
mkdir -p gpurun_out
for v in "" deep ring8; do
  lib=$PWD/simd_dct_b200/libsdct.so; [ -n "$v" ] && lib=$PWD/simd_dct_b200/variants/libsdct_$v.so
  echo "== ${v:-product}"
  SDCT_LIB=$lib timeout 300 python tools/prof_run.py --reps 20 fwd_A_tensor fwd_B_tensor fwd_C_tensor inv_A_tensor inv_B_tensor inv_C_tensor rt_B_tensor 2>&1 | tail -7
  SDCT_LIB=$lib timeout 300 python tools/prof_run.py --frames 16 --w 3840 --h 2160 --reps 10 rgb_A_tensor rgb_B_tensor rgb_C_tensor rgb_B_tensor_ycc 2>&1 | tail -4
done | tee gpurun_out/deep.txt
SDCT_LIB=$PWD/simd_dct_b200/variants/libsdct_deep.so timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_mismatch_rates.py tests/test_color.py -x -q -m gpu -k "tensor or color or mismatch" 2>&1 | tail -4
