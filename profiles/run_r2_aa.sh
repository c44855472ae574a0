#!/bin/bash
# round 2, call aa: latent record with the XU-pipe fraction; ncu --set full of the AIS kernels at HEAD (window-path k_ais_seq, k_ais_apply, k_ais_rowcnt)
mkdir -p gpurun_out/r2aa
timeout 600 python bench.py --workload latent --no-cpu-baseline --steps 10 > gpurun_out/r2aa/bench_latent.json 2> gpurun_out/r2aa/bench_latent.err; tail -2 gpurun_out/r2aa/bench_latent.err
python - <<'PY'
import json
r = json.loads([l for l in open("gpurun_out/r2aa/bench_latent.json") if l.startswith("{")][0])
print(r["value"], r["ms_per_step"], r["roofline"]["xu_pipe"]["frac"], r["roofline"]["xu_pipe"]["bound_ms_per_sweep"])
PY
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_ais_rowcnt|k_ais_seq|k_ais_apply" -s 300 -c 6 -o gpurun_out/r2aa/ais_full python profiles/time_ais.py 65536 128 > gpurun_out/r2aa/ncu_full.log 2>&1; tail -2 gpurun_out/r2aa/ncu_full.log
