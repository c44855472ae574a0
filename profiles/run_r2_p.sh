#!/bin/bash
# round 2, call p (2 GPUs): sharded parity test with the engine's own NCCL communicator, then the default 2-GPU bench line
# (replicas + the row-sharded `strong` sub-record)
mkdir -p gpurun_out/r2p
timeout 900 python -m pytest tests/test_gpu_sharded.py -q -x -m gpu --timeout 800 > gpurun_out/r2p/pytest_sharded.log 2>&1; echo "rc=$?" >> gpurun_out/r2p/pytest_sharded.log; tail -6 gpurun_out/r2p/pytest_sharded.log
( time timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --no-extras ) > gpurun_out/r2p/bench_2gpu.json 2> gpurun_out/r2p/bench_2gpu.err; echo "bench rc=$?"; tail -5 gpurun_out/r2p/bench_2gpu.err
python - <<'PY'
import json
for line in open("gpurun_out/r2p/bench_2gpu.json"):
    if line.startswith("{"):
        r = json.loads(line)
        print("weak", r["value"], r["ms_per_step"], {k: (v["value"], v["ms_per_step"]) for k, v in r["workloads"].items()})
        print("strong", {k: (v["value"], v["ms_per_step"], v["gpu_launches"]) for k, v in r.get("strong", {}).get("workloads", {}).items()})
PY
