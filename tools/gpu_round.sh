#!/bin/bash
# full GPU pass: parity tests, smoke, bench (both arms), launch list, ncu --set full of the headline and tensor-core kernels
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt; lscpu | grep -E "Model name|^CPU\(s\)|Flags" | cut -c1-400 >> gpurun_out/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.txt 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.txt
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -3 gpurun_out/bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print('headline ms', d['ms_per_step'], 'GB/s', d['gbps'], 'frac', d['roofline']['frac'])
print('e2e', d['e2e']['ms_per_step'], d['e2e']['gbps_each_way'])
for k,v in d['variants'].items(): print(k, round(v['ms_per_step'],4), round(v['gbps_per_gpu']), round(v['frac_of_hbm_peak'],3))
print(d['clocks']); print(d.get('cpu_baseline'))
PY
timeout 300 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err; echo "ref rc=$?"
timeout 300 python bench.py --steps 3 --warmup 3 --no-variants > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-variants > gpurun_out/ncu.log 2>&1
echo "ncu launches rc=$?"
bash tools/gpu_ncu.sh fwd_B_exact forward_kernel --reps 4 fwd_B_exact
bash tools/gpu_ncu.sh fwd_B_tensor forward_tc_kernel --reps 4 fwd_B_tensor
bash tools/gpu_ncu.sh fwd_A_tensor forward_tc_kernel --reps 4 fwd_A_tensor
bash tools/gpu_ncu.sh inv_B inverse_kernel --reps 4 inv_B
bash tools/gpu_ncu.sh inv_B_tensor inverse_tc_kernel --reps 4 inv_B_tensor
timeout 300 python tools/prof_run.py --frames 16 --w 3840 --h 2160 --reps 10 rgb_B_exact rgb_B_fast rgb_B_tensor rgb_A_tensor rgb_C_tensor fwd_B_tensor fwd_A_tensor fwd_C_tensor fwd_B_exact inv_A inv_B inv_C inv_A_tensor inv_B_tensor inv_C_tensor > gpurun_out/time_4k.txt 2>&1; cat gpurun_out/time_4k.txt
