#!/bin/bash
# short pass: bench (both arms), 4K timings, ncu --set full of the layout-C tensor-core kernels
mkdir -p gpurun_out
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print('headline ms', d['ms_per_step'], 'GB/s', d['gbps'], 'frac', d['roofline']['frac'])
print('e2e', d['e2e']['ms_per_step'], d['e2e']['gbps_each_way'])
for k,v in d['variants'].items(): print(k, round(v['ms_per_step'],4), round(v['gbps_per_gpu']), round(v['frac_of_hbm_peak'],3))
print(d['clocks']); print(d.get('cpu_baseline'))
PY
timeout 300 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err; echo "ref rc=$?"
bash tools/gpu_ncu.sh fwd_C_tensor forward_tc_kernel --reps 4 fwd_C_tensor
bash tools/gpu_ncu.sh inv_C_tensor inverse_tc_kernel --reps 4 inv_C_tensor
bash tools/gpu_ncu.sh inv_A_tensor inverse_tc_kernel --reps 4 inv_A_tensor
bash tools/gpu_ncu.sh inv_B_tensor inverse_tc_kernel --reps 4 inv_B_tensor
timeout 300 python tools/prof_run.py --frames 16 --w 3840 --h 2160 --reps 10 rgb_B_exact rgb_B_fast rgb_B_tensor rgb_A_tensor rgb_C_tensor fwd_B_tensor fwd_A_tensor fwd_C_tensor fwd_B_exact inv_A inv_B inv_C inv_A_tensor inv_B_tensor inv_C_tensor > gpurun_out/time_4k.txt 2>&1; cat gpurun_out/time_4k.txt
