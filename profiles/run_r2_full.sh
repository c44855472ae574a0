#!/bin/bash
# full validation pass at HEAD on one B200: every -m gpu test, smoke, the default bench line
mkdir -p gpurun_out/r2v; O=gpurun_out/r2v
timeout 1500 python -m pytest tests -q -m gpu --durations=10 > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" > $O/rc.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/rc.txt
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -16 $O/pytest_gpu.log; tail -2 $O/smoke.log
