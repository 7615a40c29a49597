#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout 300 python tools/prof_run.py --reps 20 fwd_A_tensor fwd_B_tensor fwd_C_tensor inv_A_tensor inv_B_tensor inv_C_tensor rt_A_tensor rt_B_tensor rt_C_tensor 2>&1 | tail -9 | tee gpurun_out/ring8.txt
timeout 300 python tools/prof_run.py --frames 16 --w 3840 --h 2160 --reps 10 fwd_A_tensor fwd_B_tensor fwd_C_tensor inv_A_tensor inv_B_tensor inv_C_tensor rgb_A_tensor rgb_B_tensor rgb_C_tensor irgb_A_tensor irgb_B_tensor irgb_C_tensor 2>&1 | tail -12 | tee -a gpurun_out/ring8.txt
