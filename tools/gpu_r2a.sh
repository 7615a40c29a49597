#!/bin/bash
# round-2 pass A: parity tests, conversion variants side by side, bench (both arms), ncu --set full of the EXACT kernel
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt; lscpu | grep -E "Model name|^CPU\(s\)|NUMA" >> gpurun_out/nproc.txt
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.txt 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.txt
for v in "" $VARIANTS; do
  if [ -n "$v" ]; then export SDCT_LIB=$PWD/simd_dct_b200/variants/libsdct_$v.so; else unset SDCT_LIB; fi
  echo "== variant '${v:-default}'"
  timeout 300 python tools/prof_run.py --reps 30 fwd_B_exact fwd_A_exact fwd_C_exact fwd_A_fast fwd_B_fast fwd_C_fast rt_B_exact 2>&1 | tail -8
done
unset SDCT_LIB
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -5 gpurun_out/bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print('headline ms', d['ms_per_step'], 'GB/s', d['gbps'], 'frac', d['roofline']['frac'])
print('parity', d['parity'])
e=d['e2e']; print('e2e', e['ms_per_step'], e['gbps_each_way'], 'ceiling', e['copy_ceiling'], 'frac', e['frac_of_copy_ceiling']); print(e.get('other_entry_points'))
for c,vs in d['variants'].items():
    print('##', c)
    for k,v in vs.items():
        if isinstance(v, dict) and 'us_per_launch' in v: print('  ', k, round(v['us_per_launch'],1), 'us', round(v['gbps_per_gpu']), 'GB/s', round(v['frac_of_hbm_peak'],3))
        else: print('  ', k, v)
print(d['clocks']); print(json.dumps(d.get('cpu_baseline'), indent=1))
PY
timeout 300 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json | cut -c1-400
bash tools/gpu_ncu.sh fwd_B_exact forward_kernel --reps 4 fwd_B_exact
