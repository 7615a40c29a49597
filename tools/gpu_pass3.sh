#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.txt 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_gpu.txt
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -5 gpurun_out/bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print('headline ms', d['ms_per_step'], 'GB/s', d['gbps'], 'frac', d['roofline']['frac'])
print('e2e', d['e2e']['ms_per_step'], d['e2e']['gbps_each_way'])
for k,v in d['variants'].items(): print(k, round(v['ms_per_step'],4), round(v['gbps_per_gpu']), round(v['frac_of_hbm_peak'],3))
print(d['clocks'])
PY
