#!/bin/bash
mkdir -p gpurun_out/r2c; O=gpurun_out/r2c
timeout 1500 python -m pytest tests/test_gpu_posterior_moments.py tests/test_gpu_mf_probit.py tests/test_gpu_weighted_models.py tests/test_gpu_sharded.py -q -m gpu --durations=8 > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" > $O/rc.txt
timeout 300 python bench.py --workload eigen --steps 10 > $O/bench_eigen.json 2> $O/bench_eigen.err; echo "bench eigen rc=$?" >> $O/rc.txt
( time timeout 900 python bench.py --impl reference ) > $O/bench_ref.json 2> $O/bench_ref.err; echo "bench ref rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -30 $O/pytest_gpu.log; cut -c1-700 $O/bench_eigen.json; tail -5 $O/bench_ref.err; cut -c1-2500 $O/bench_ref.json
