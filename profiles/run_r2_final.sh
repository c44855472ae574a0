#!/bin/bash
# round 2, last call: every -m gpu test at HEAD, then smoke
mkdir -p gpurun_out/r2final
( time timeout 1500 python -m pytest tests -q -x -m gpu --timeout 1200 ) > gpurun_out/r2final/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/r2final/pytest_gpu.log; tail -6 gpurun_out/r2final/pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2final/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/r2final/smoke.log; tail -2 gpurun_out/r2final/smoke.log
