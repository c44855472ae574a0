#!/bin/bash
# round 2, call t: full validation at HEAD on one B200: every -m gpu test, smoke, the default bench line, ncu launch list of it
mkdir -p gpurun_out/r2t
( time timeout 2400 python -m pytest tests -q -x -m gpu --timeout 1500 ) > gpurun_out/r2t/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/r2t/pytest_gpu.log; tail -6 gpurun_out/r2t/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2t/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/r2t/smoke.log; tail -2 gpurun_out/r2t/smoke.log
( time timeout 900 python bench.py ) > gpurun_out/r2t/bench_default.json 2> gpurun_out/r2t/bench_default.err; tail -4 gpurun_out/r2t/bench_default.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2t/launches_sbm.csv python bench.py --workload sbm --steps 3 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r2t/ncu_sbm.log 2>&1; tail -1 gpurun_out/r2t/ncu_sbm.log | cut -c1-200
