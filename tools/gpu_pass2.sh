#!/bin/bash
# second GPU pass: packed-fp32 probe, fixed bench, launch list + full ncu capture of the headline kernel
mkdir -p gpurun_out
./tools/probe2 > gpurun_out/probe2.txt 2>&1; echo "probe2 rc=$?"
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
head -c 1500 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
for mb in 2 4 8; do SDCT_HOST_CHUNK_MB=$mb timeout 300 python bench.py --steps 10 --warmup 3 --no-variants --preheat 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('chunk_mb', $mb, d['e2e']['ms_per_step'], d['e2e']['gbps_each_way'])"; done > gpurun_out/e2e_chunks.txt 2>&1
cat gpurun_out/e2e_chunks.txt
timeout 300 python bench.py --steps 3 --warmup 3 --no-variants --no-e2e --preheat 0 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:forward_kernel -s 3 -c 2 -f -o gpurun_out/fwdB_exact python bench.py --steps 3 --warmup 3 --no-variants --no-e2e --preheat 0 > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"
