#!/bin/bash
mkdir -p gpurun_out/r2af
timeout 100 python -c "import __graft_entry__ as g; g.smoke(); print('smoke OK')" > gpurun_out/r2af/smoke.log 2>&1; tail -1 gpurun_out/r2af/smoke.log
timeout 120 python -m pytest tests/test_gpu_sbm.py tests/test_gpu_ais_batched.py tests/test_gpu_eigen.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r2af/pytest.log 2>&1; tail -2 gpurun_out/r2af/pytest.log
