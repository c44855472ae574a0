#!/bin/bash
# round 2, call r (8 GPUs): sharded parity at 2 / 4 / 8 ranks with the engine's own communicator, then the default 8-GPU bench
# line (replicas + row-sharded strong sub-record + AIS particles partitioned)
mkdir -p gpurun_out/r2r
timeout 900 python -m pytest tests/test_gpu_sharded.py -q -m gpu --timeout 800 > gpurun_out/r2r/pytest_sharded.log 2>&1; echo "rc=$?" >> gpurun_out/r2r/pytest_sharded.log; tail -6 gpurun_out/r2r/pytest_sharded.log
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --steps 20 --warmup 3 ) > gpurun_out/r2r/bench_8gpu.json 2> gpurun_out/r2r/bench_8gpu.err; echo "bench rc=$?"; tail -5 gpurun_out/r2r/bench_8gpu.err
python - <<'PY'
import json
for line in open("gpurun_out/r2r/bench_8gpu.json"):
    if line.startswith("{"):
        r = json.loads(line)
        print("weak", r["value"], r["ms_per_step"], r.get("n_changed_last_sweep"), r.get("extra_warmup_sweeps"), {k: (round(v["value"],1), round(v["ms_per_step"],2)) for k, v in r["workloads"].items()})
        print("strong", {k: (round(v["value"],1), round(v["ms_per_step"],2), v["gpu_launches"]) for k, v in r.get("strong", {}).get("workloads", {}).items()})
PY
