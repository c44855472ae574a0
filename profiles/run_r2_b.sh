#!/bin/bash
# Round-2 second pass: new GPU tests (config sizes, posterior moments, probit MF, weighted models), smoke, the new one-line bench.
mkdir -p gpurun_out/r2b; O=gpurun_out/r2b
nproc > $O/nproc.txt; free -g >> $O/nproc.txt
timeout 1500 python -m pytest tests -q -m gpu -x --durations=15 > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" > $O/rc.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/rc.txt
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -25 $O/pytest_gpu.log; tail -3 $O/smoke.log; tail -5 $O/bench.err; cut -c1-3000 $O/bench.json
