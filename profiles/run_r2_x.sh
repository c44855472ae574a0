#!/bin/bash
# round 2, call x: validation at HEAD on one B200: every -m gpu test, smoke, default bench line, launch list of the SBM + AIS sweeps,
# ncu --set full of the rewritten small SBM kernels
mkdir -p gpurun_out/r2x
( time timeout 2400 python -m pytest tests -q -x -m gpu --timeout 1500 ) > gpurun_out/r2x/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/r2x/pytest_gpu.log; tail -6 gpurun_out/r2x/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2x/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/r2x/smoke.log; tail -2 gpurun_out/r2x/smoke.log
( time timeout 900 python bench.py ) > gpurun_out/r2x/bench_default.json 2> gpurun_out/r2x/bench_default.err; tail -4 gpurun_out/r2x/bench_default.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 1500 -c 600 --csv --log-file gpurun_out/r2x/launches_sbm.csv python bench.py --workload sbm --steps 6 --warmup 6 --no-cpu-baseline --no-extras > gpurun_out/r2x/ncu_sbm.log 2>&1; tail -1 gpurun_out/r2x/ncu_sbm.log | cut -c1-120
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_lp_from_stats|k_stats_fix|k_strip_stats|k_sbm_seq2" -s 200 -c 8 -o gpurun_out/r2x/sbm_full python bench.py --workload sbm --steps 6 --warmup 6 --no-cpu-baseline --no-extras > gpurun_out/r2x/ncu_sbm_full.log 2>&1; tail -1 gpurun_out/r2x/ncu_sbm_full.log | cut -c1-120
