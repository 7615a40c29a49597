#!/bin/bash
# tensor-core kernel bring-up: parity test (short timeout: a deadlocked role chain must not eat the budget), then timing
mkdir -p gpurun_out
timeout -k 5 ${TO:-90} python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tensor_core" > gpurun_out/tc_pytest.txt 2>&1; rc=$?; echo "pytest rc=$rc"; tail -30 gpurun_out/tc_pytest.txt
[ $rc -ne 0 ] && exit 1
timeout -k 5 ${TO:-90} python tools/prof_run.py --reps 20 fwd_B_tensor fwd_A_tensor fwd_C_tensor fwd_B_fast fwd_B_exact > gpurun_out/tc_time.txt 2>&1; echo "time rc=$?"; cat gpurun_out/tc_time.txt
