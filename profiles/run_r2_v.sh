#!/bin/bash
mkdir -p gpurun_out/r2v
timeout 1500 python -m pytest tests/test_gpu_sbm.py -q -x -m gpu --timeout 900 > gpurun_out/r2v/pytest_sbm.log 2>&1; tail -3 gpurun_out/r2v/pytest_sbm.log
