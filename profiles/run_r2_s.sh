#!/bin/bash
# round 2, call s: SBM tests after the shared-memory k_lp_from_stats, SBM bench, collapsed sampler test
mkdir -p gpurun_out/r2s
timeout 1500 python -m pytest tests/test_gpu_sbm.py tests/test_gpu_model_sbm.py tests/test_gpu_newapi.py tests/test_gpu_ais.py tests/test_gpu_config_sizes.py -q -x -m gpu --timeout 900 > gpurun_out/r2s/pytest_sbm.log 2>&1; echo "rc=$?" >> gpurun_out/r2s/pytest_sbm.log; tail -4 gpurun_out/r2s/pytest_sbm.log
timeout 600 python bench.py --workload sbm --no-cpu-baseline > gpurun_out/r2s/bench_sbm.json 2> gpurun_out/r2s/bench_sbm.err; tail -2 gpurun_out/r2s/bench_sbm.err
python - <<'PY'
import json
r = json.loads([l for l in open("gpurun_out/r2s/bench_sbm.json") if l.startswith("{")][0])
print(r["value"], r["ms_per_step"], r["e2e"]["value"], r["roofline"]["kernel_ms_per_sweep"], r["roofline"]["other_kernels_ms_per_sweep"], r["mixing"]["first_sweep_from_prior_labels_ms"])
PY
