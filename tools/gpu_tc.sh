#!/bin/bash
# tensor-core kernel bring-up: parity test, then timing of the three layouts
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tensor_core" > gpurun_out/tc_pytest.txt 2>&1; echo "pytest rc=$?"; tail -30 gpurun_out/tc_pytest.txt
timeout 300 python tools/prof_run.py --reps 20 fwd_B_tensor fwd_A_tensor fwd_C_tensor fwd_B_fast fwd_B_exact > gpurun_out/tc_time.txt 2>&1; echo "time rc=$?"; cat gpurun_out/tc_time.txt
