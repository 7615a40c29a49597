#!/bin/bash
# quick pass: parity tests (pytest -k "$K", default: the forward ones) + timings of $NAMES, once per entry of $ENVS
# (each entry = "VAR=value,VAR2=value" exported for that run; "-" = nothing) and per experiment build in $VARIANTS
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q ${K:+-k "$K"} > gpurun_out/pytest_quick.txt 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_quick.txt
for v in "" $VARIANTS; do
  if [ -n "$v" ]; then export SDCT_LIB=$PWD/simd_dct_b200/variants/libsdct_$v.so; else unset SDCT_LIB; fi
  for e in ${ENVS:--}; do
    echo "== variant '${v:-default}' env '$e'"
    ( if [ "$e" != "-" ]; then IFS=, ; for kv in $e; do export "$kv"; done; unset IFS; fi
      timeout 300 python tools/prof_run.py --reps ${REPS:-30} $PROF_ARGS ${NAMES:-fwd_B_exact fwd_A_exact fwd_C_exact fwd_A_fast fwd_B_fast fwd_C_fast} 2>&1 | tail -12 )
  done
done
