#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
for v in "" rgb5 rt6; do
  lib=$PWD/simd_dct_b200/libsdct.so; [ -n "$v" ] && lib=$PWD/simd_dct_b200/variants/libsdct_$v.so
  echo "== ${v:-product}"
  SDCT_LIB=$lib timeout 300 python tools/prof_run.py --reps 20 rt_A_tensor rt_B_tensor rt_C_tensor rt_B_tensor 2>&1 | tail -4
  SDCT_LIB=$lib timeout 300 python tools/prof_run.py --frames 16 --w 3840 --h 2160 --reps 10 rgb_A_tensor rgb_B_tensor rgb_C_tensor irgb_A_tensor irgb_B_tensor irgb_C_tensor rgb_B_tensor 2>&1 | tail -7
done | tee gpurun_out/rings2.txt
