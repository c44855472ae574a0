#!/bin/bash
# round 2, call ab: SBM bench with whole-sweep speculation (mode 3) against the block-by-block overlap (mode 1)
mkdir -p gpurun_out/r2ab
for mode in 3 1; do
GK_SBM_OVERLAP_MODE=$mode timeout 600 python bench.py --workload sbm --no-cpu-baseline --steps 20 > gpurun_out/r2ab/bench_sbm_mode$mode.json 2> gpurun_out/r2ab/bench_sbm_mode$mode.err; tail -1 gpurun_out/r2ab/bench_sbm_mode$mode.err
done
python - <<'PY'
import json
for mode in (3, 1):
    r = json.loads([l for l in open("gpurun_out/r2ab/bench_sbm_mode%d.json" % mode) if l.startswith("{")][0])
    m = r["mixing"]
    print(mode, round(r["value"],1), round(r["ms_per_step"],3), "e2e", round(r["e2e"]["value"],1), "strip", round(r["roofline"]["kernel_ms_per_sweep"],3), r["roofline"]["launches_per_sweep"], round(r["roofline"]["frac"],3), r["roofline"]["other_kernels_ms_per_sweep"], "mix", round(m["first_sweep_from_prior_labels_ms"],1), round(m["weak_signal"]["ms_per_step"],1), round(m["no_signal"]["ms_per_step"],1), "f64", round(r["workloads"]["sbm_f64"]["ms_per_step"],3), r["n_changed_last_sweep"])
PY
