#!/bin/bash
# quick pass: forward parity tests + timings of the fp32 kernels (optionally for $VARIANTS)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "config1 or ragged or band_window or unaligned or interleaved or fast or batch or full_size or dct8" > gpurun_out/pytest_quick.txt 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_quick.txt
for v in "" $VARIANTS; do
  if [ -n "$v" ]; then export SDCT_LIB=$PWD/simd_dct_b200/variants/libsdct_$v.so; else unset SDCT_LIB; fi
  echo "== variant '${v:-default}'"
  timeout 300 python tools/prof_run.py --reps 30 ${NAMES:-fwd_B_exact fwd_A_exact fwd_C_exact fwd_A_fast fwd_B_fast fwd_C_fast} 2>&1 | tail -8
done
