#!/bin/bash
# profiling: the host call's per-chunk timeline ($SDCT_HOST_TRACE) and a sweep of the pipeline's knobs (one GPU)
mkdir -p gpurun_out
SDCT_HOST_TRACE=1 timeout 120 python tools/e2e_probe.py > gpurun_out/e2e_trace.txt 2>&1
tail -40 gpurun_out/e2e_trace.txt
for mb in 2 4 8 16 32; do for slots in 3 4 8; do
  echo -n "slots=$slots "; SDCT_HOST_SLOTS=$slots SDCT_HOST_CHUNK_MB=$mb timeout 120 python tools/e2e_probe.py 2>&1 | tail -1
done; done | tee gpurun_out/e2e_sweep.txt
echo -n "no ramp "; SDCT_HOST_NO_RAMP=1 timeout 120 python tools/e2e_probe.py 2>&1 | tail -1 | tee -a gpurun_out/e2e_sweep.txt
timeout 120 python tools/pcie_probe.py --gpus 1 --chunk-mb 16 2>&1 | tail -3 | tee -a gpurun_out/e2e_sweep.txt
timeout 120 python tools/pcie_probe.py --gpus 1 2>&1 | tail -3 | tee -a gpurun_out/e2e_sweep.txt
