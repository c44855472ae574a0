#!/bin/bash
# round 2, call w (8 GPUs): the default bench line at HEAD as the driver launches it (replicas on identical data, strong sub-record, AIS)
mkdir -p gpurun_out/r2w
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 8 --steps 20 --warmup 3 ) > gpurun_out/r2w/bench_8gpu.json 2> gpurun_out/r2w/bench_8gpu.err; echo "bench rc=$?"; tail -4 gpurun_out/r2w/bench_8gpu.err
python - <<'PY'
import json
for line in open("gpurun_out/r2w/bench_8gpu.json"):
    if line.startswith("{"):
        r = json.loads(line)
        print("weak", r["value"], r["ms_per_step"], r.get("n_changed_last_sweep"), r.get("extra_warmup_sweeps"), r["e2e"]["value"], {k: (round(v["value"],1), round(v["ms_per_step"],2)) for k, v in r["workloads"].items()})
        print("strong", {k: (round(v["value"],1), round(v["ms_per_step"],2), v["gpu_launches"]) for k, v in r.get("strong", {}).get("workloads", {}).items()})
        print(r["parallelism"], r["clocks"])
PY
