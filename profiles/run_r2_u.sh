#!/bin/bash
# round 2, call u: SBM tests + SBM bench with the mixing records
mkdir -p gpurun_out/r2u
timeout 1500 python -m pytest tests/test_gpu_sbm.py tests/test_gpu_model_sbm.py tests/test_gpu_config_sizes.py -q -x -m gpu --timeout 900 > gpurun_out/r2u/pytest_sbm.log 2>&1; tail -2 gpurun_out/r2u/pytest_sbm.log
timeout 600 python bench.py --workload sbm --no-cpu-baseline --steps 10 > gpurun_out/r2u/bench_sbm.json 2> gpurun_out/r2u/bench_sbm.err; tail -2 gpurun_out/r2u/bench_sbm.err
GK_SBM_NO_OVERLAP=1 timeout 600 python bench.py --workload sbm --no-cpu-baseline --steps 10 > gpurun_out/r2u/bench_sbm_nooverlap.json 2> gpurun_out/r2u/bench_sbm_nooverlap.err
python - <<'PY'
import json
for f in ("bench_sbm", "bench_sbm_nooverlap"):
    r = json.loads([l for l in open("gpurun_out/r2u/%s.json" % f) if l.startswith("{")][0])
    m = r["mixing"]
    print(f, round(r["value"],1), round(r["ms_per_step"],3), "first", round(m["first_sweep_from_prior_labels_ms"],1), "weak", round(m["weak_signal"]["ms_per_step"],1), m["weak_signal"]["n_changed_per_sweep"], "none", round(m["no_signal"]["ms_per_step"],1), round(m["no_signal"]["us_per_move"],2))
PY
