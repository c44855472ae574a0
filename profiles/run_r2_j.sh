#!/bin/bash
# round 2, call j: static-scratch entry point alone in a fresh process, then after other GPU tests (it is collected last)
mkdir -p gpurun_out/r2j
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -x > gpurun_out/r2j/alone.log 2>&1; echo "rc=$?" >> gpurun_out/r2j/alone.log
timeout 1500 python -m pytest tests/test_gpu_probit.py tests/test_gpu_eigen.py tests/test_gpu_sbm.py tests/test_zz_no_alloc.py -q -x -m gpu --timeout 900 > gpurun_out/r2j/after_others.log 2>&1; echo "rc=$?" >> gpurun_out/r2j/after_others.log
tail -3 gpurun_out/r2j/alone.log; tail -4 gpurun_out/r2j/after_others.log
