#!/bin/bash
mkdir -p gpurun_out/r2ad
timeout 500 python bench.py --workload ais --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/r2ad/bench_ais.json 2> gpurun_out/r2ad/bench_ais.err; tail -2 gpurun_out/r2ad/bench_ais.err
python - <<'PY'
import json
r = json.loads([l for l in open("gpurun_out/r2ad/bench_ais.json") if l.startswith("{")][0])
print(r["value"], r["stationary"]["value"], r["ladder"])
PY
