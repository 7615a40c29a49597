#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_color.py -x -q -m gpu -k "tensor or color" 2>&1 | tail -3
timeout 300 python tools/prof_run.py --reps 20 fwd_A_tensor fwd_B_tensor fwd_C_tensor fwd_A_tensor fwd_B_tensor fwd_C_tensor 2>&1 | tail -6 | tee gpurun_out/addr.txt
timeout 300 python tools/prof_run.py --frames 16 --w 3840 --h 2160 --reps 10 rgb_A_tensor rgb_B_tensor rgb_C_tensor rgb_B_tensor_ycc 2>&1 | tail -4 | tee -a gpurun_out/addr.txt
