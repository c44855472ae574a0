#!/bin/bash
# round 2, call k: batched AIS (gk_ais_*): parity tests, then timing at N = 65536
mkdir -p gpurun_out/r2k
timeout 900 python -m pytest tests/test_gpu_ais_batched.py -q -x -m gpu --timeout 600 > gpurun_out/r2k/pytest_ais.log 2>&1; echo "rc=$?" >> gpurun_out/r2k/pytest_ais.log
tail -15 gpurun_out/r2k/pytest_ais.log
timeout 600 python profiles/time_ais.py 16384 8 128 > gpurun_out/r2k/time_ais_16k.log 2>&1; tail -6 gpurun_out/r2k/time_ais_16k.log
timeout 900 python profiles/time_ais.py 65536 8 32 128 > gpurun_out/r2k/time_ais_64k.log 2>&1; tail -8 gpurun_out/r2k/time_ais_64k.log
