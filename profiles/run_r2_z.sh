#!/bin/bash
# round 2, call z (2 GPUs): default bench line at HEAD under torchrun (replicas + strong + AIS with incremental counts), then the full -m gpu suite (the sharded test runs at 2 ranks)
mkdir -p gpurun_out/r2z
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29515 bench.py --gpus 2 --steps 20 --warmup 3 ) > gpurun_out/r2z/bench_2gpu.json 2> gpurun_out/r2z/bench_2gpu.err; echo "bench rc=$?"; tail -3 gpurun_out/r2z/bench_2gpu.err
python - <<'PY'
import json
for line in open("gpurun_out/r2z/bench_2gpu.json"):
    if line.startswith("{"):
        r = json.loads(line)
        print("weak", round(r["value"],1), round(r["ms_per_step"],3), r.get("n_changed_last_sweep"), {k: (round(v["value"],1), round(v["ms_per_step"],2)) for k, v in r["workloads"].items()})
        print("ais stationary", r["workloads"]["ais"]["stationary"])
        print("strong", {k: (round(v["value"],1), round(v["ms_per_step"],2)) for k, v in r.get("strong", {}).get("workloads", {}).items()})
PY
( time timeout 2400 python -m pytest tests -q -x -m gpu --timeout 1500 ) > gpurun_out/r2z/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/r2z/pytest_gpu.log; tail -5 gpurun_out/r2z/pytest_gpu.log
