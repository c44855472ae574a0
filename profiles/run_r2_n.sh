#!/bin/bash
# round 2, call n: AIS after the cp.async ring: parity tests, bench record
mkdir -p gpurun_out/r2n
timeout 900 python -m pytest tests/test_gpu_ais_batched.py tests/test_gpu_ais.py -q -x -m gpu --timeout 600 > gpurun_out/r2n/pytest_ais.log 2>&1; echo "rc=$?" >> gpurun_out/r2n/pytest_ais.log
tail -5 gpurun_out/r2n/pytest_ais.log
timeout 600 python bench.py --workload ais --steps 6 --warmup 3 > gpurun_out/r2n/bench_ais_p128.json 2> gpurun_out/r2n/bench_ais_p128.err; tail -3 gpurun_out/r2n/bench_ais_p128.err
python - <<'PY'
import json
r = json.load(open("gpurun_out/r2n/bench_ais_p128.json"))
print(r["value"], r["ms_per_step"], r["e2e"]["value"], r["particle_scan"], r["roofline"]["kernel_ms_per_sweep"], r["roofline"]["other_kernels_ms_per_sweep"])
PY
timeout 600 python profiles/time_ais.py 65536 128 > gpurun_out/r2n/time_ais_64k.log 2>&1; tail -2 gpurun_out/r2n/time_ais_64k.log
