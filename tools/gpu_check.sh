#!/bin/bash
# quick GPU pass: parity tests, smoke, bench, tensor-core timing and role trace
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.txt 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.txt 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.txt
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json | cut -c1-3000
timeout 300 python tools/prof_run.py --reps 20 fwd_B_tensor fwd_A_tensor fwd_C_tensor fwd_B_fast fwd_B_exact > gpurun_out/tc_time.txt 2>&1; echo "time rc=$?"; cat gpurun_out/tc_time.txt
ABL="0" LAYOUT=B bash tools/gpu_trace.sh | head -80
