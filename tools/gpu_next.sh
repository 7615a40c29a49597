#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_host_path.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -4
for m in 0 1; do echo -n "engine_streams=$m "; SDCT_HOST_ENGINE_STREAMS=$m timeout 120 python tools/e2e_probe.py 2>&1 | tail -1; done
