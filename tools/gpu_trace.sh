#!/bin/bash
mkdir -p gpurun_out
for a in ${ABL:-31 0}; do timeout 120 python tools/tc_trace.py $a ${LAYOUT:-A}; done > gpurun_out/trace.txt 2>&1
cat gpurun_out/trace.txt
