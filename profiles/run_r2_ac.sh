#!/bin/bash
# round 2, call ac (2 GPUs): default bench line at HEAD (speculated strip passes on) under torchrun
mkdir -p gpurun_out/r2ac
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29516 bench.py --gpus 2 --steps 20 --warmup 3 ) > gpurun_out/r2ac/bench_2gpu.json 2> gpurun_out/r2ac/bench_2gpu.err; echo "bench rc=$?"; tail -3 gpurun_out/r2ac/bench_2gpu.err
python - <<'PY'
import json
for line in open("gpurun_out/r2ac/bench_2gpu.json"):
    if line.startswith("{"):
        r = json.loads(line)
        print("weak", round(r["value"],1), round(r["ms_per_step"],3), r.get("n_changed_last_sweep"), {k: (round(v["value"],1), round(v["ms_per_step"],2)) for k, v in r["workloads"].items()})
        print("strong", {k: (round(v["value"],1), round(v["ms_per_step"],2)) for k, v in r.get("strong", {}).get("workloads", {}).items()})
PY
