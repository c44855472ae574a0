#!/bin/bash
mkdir -p gpurun_out/r2ae
timeout 400 python bench.py > gpurun_out/r2ae/bench_default_1gpu.json 2> gpurun_out/r2ae/bench_default.err; echo rc=$?; tail -2 gpurun_out/r2ae/bench_default.err
python - <<'PY'
import json
r = json.loads([l for l in open("gpurun_out/r2ae/bench_default_1gpu.json") if l.startswith("{")][0])
print(r["value"], r["ms_per_step"], r["e2e"]["value"], r["roofline"]["frac"], r["gpu_launches"], r["clocks"])
for k, w in r["workloads"].items(): print(k, w.get("value"), w.get("ms_per_step"))
print(r["workloads"]["ais"].get("ladder"))
PY
