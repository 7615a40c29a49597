#!/usr/bin/env python
"""bench_summary.py <bench.json> - the bench line as a readable table (stderr-free; used by tools/gpu_round.sh)"""
import json
import sys

d = json.load(open(sys.argv[1]))
print("headline: %.1f us/step, %.3g blocks/s, %.0f GB/s, frac %.3f of %s" % (d["ms_per_step"] * 1e3, d["value"], d["gbps"], d["roofline"]["frac"], d["roofline"]["peak"]))
print("parity:", d.get("parity"))
e = d.get("e2e")
if e:
    print("e2e: %.3f ms/step, %.3g blocks/s, %.1f GB/s each way; copy ceiling %.1f GB/s/GPU (%.1f aggregate) -> frac %.2f" % (
        e["ms_per_step"], e["value"], e["gbps_each_way"], e["copy_ceiling"]["gbps_each_way_per_gpu"], e["copy_ceiling"]["aggregate_gbps_each_way"], e["frac_of_copy_ceiling"]))
    for k, v in (e.get("other_entry_points") or {}).items():
        print("   ", k, "%.3f ms, %.1f GB/s each way" % (v["ms_per_step"], v["gbps_each_way"]))
    if e.get("single_process_all_gpus"):
        print("   ", "single process:", e["single_process_all_gpus"])
for c, vs in (d.get("variants") or {}).items():
    print("##", c)
    for k, v in vs.items():
        if isinstance(v, dict) and "us_per_launch" in v:
            print("   %-52s %9.1f us %6.0f GB/s  %.3f" % (k, v["us_per_launch"], v["gbps_per_gpu"], v["frac_of_hbm_peak"]))
        else:
            print("   %-52s %s" % (k, v))
print("clocks:", d.get("clocks"))
cb = d.get("cpu_baseline") or {}
print("cpu: %.3g blocks/s on %s cores (%s)" % (cb.get("value", 0), cb.get("cores"), cb.get("sample")))
for k, v in cb.items():
    if isinstance(v, dict):
        print("   ", k, "%.3g blocks/s, %s" % (v["value"], v["sample"]))
