#!/bin/bash
# full GPU evidence pass of a round (one GPU): parity tests, smoke, bench (both arms), launch list, ncu --set full of the
# headline and the other dominant kernels, 4K timing table.  Everything lands in gpurun_out/; copy what is to be judged
# into profiles/ (tools/ncu_summary.py turns a .ncu-rep into the csv summaries kept there).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt; lscpu | grep -E "Model name|^CPU\(s\)|NUMA" >> gpurun_out/nproc.txt
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.txt 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.txt
timeout 300 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -3 gpurun_out/bench.err
python tools/bench_summary.py gpurun_out/bench.json
timeout 300 python bench.py --steps 3 --warmup 3 --no-variants --no-e2e > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-variants --no-e2e > gpurun_out/ncu.log 2>&1
echo "ncu launches rc=$?"
for spec in "fwd_B_exact forward_kernel" "fwd_A_exact forward_kernel" "rt_B_exact roundtrip_kernel" "inv_B_exact inverse_kernel" "fwd_B_tensor forward_tc_kernel" \
            "fwd_A_tensor forward_tc_kernel" "fwd_C_tensor forward_tc_kernel" "inv_B_tensor inverse_tc_kernel" "inv_A_tensor inverse_tc_kernel" "inv_C_tensor inverse_tc_kernel" "rt_B_tensor roundtrip_tc_kernel" $EXTRA_NCU; do
  set -- $spec
  timeout 300 bash tools/gpu_ncu.sh $1 $2 --reps 4 $1 2>&1 | tail -2
done
timeout 300 bash tools/gpu_ncu.sh rgb_B_tensor forward_tc_kernel --frames 16 --w 3840 --h 2160 --reps 4 rgb_B_tensor 2>&1 | tail -2
timeout 300 bash tools/gpu_ncu.sh irgb_B_tensor inverse_tc_kernel --frames 16 --w 3840 --h 2160 --reps 4 irgb_B_tensor 2>&1 | tail -2
timeout 300 python tools/prof_run.py --frames 16 --w 3840 --h 2160 --reps 10 fwd_A_exact fwd_B_exact fwd_C_exact fwd_A_fast fwd_B_fast fwd_C_fast fwd_A_tensor fwd_B_tensor fwd_C_tensor \
  inv_A_exact inv_B_exact inv_C_exact inv_A_tensor inv_B_tensor inv_C_tensor rt_A_exact rt_B_exact rt_C_exact rt_A_tensor rt_B_tensor rt_C_tensor \
  rgb_A_tensor rgb_B_tensor rgb_C_tensor rgb_B_exact rgb_B_fast rgb_B_tensor_ycc irgb_A_tensor irgb_B_tensor irgb_C_tensor irgb_B_exact irgb_B_tensor_ycc rtrgb_B_exact rtrgb_B_tensor > gpurun_out/time_4k.txt 2>&1
cat gpurun_out/time_4k.txt
