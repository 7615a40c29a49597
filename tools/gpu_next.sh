#!/bin/bash
# round-2 follow-up pass (one GPU): role stamps of the three tensor-core kernels (trace build), host pipeline timeline + knob sweep
mkdir -p gpurun_out
T=$PWD/simd_dct_b200/variants/libsdct_trace.so
if [ -f $T ]; then
  for spec in "B roundtrip" "B inverse" "B" "C"; do
    SDCT_LIB=$T timeout 120 python tools/tc_trace.py 0 $spec
  done > gpurun_out/tc_traces.txt 2>&1
  tail -30 gpurun_out/tc_traces.txt
fi
bash tools/gpu_e2e.sh
