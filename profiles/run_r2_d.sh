#!/bin/bash
mkdir -p gpurun_out/r2d; O=gpurun_out/r2d
timeout 600 python -m pytest tests/test_gpu_eigen.py tests/test_gpu_probit.py tests/test_gpu_geweke.py tests/test_gpu_config_sizes.py -q -m gpu -x > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" > $O/rc.txt
timeout 300 python bench.py --workload eigen --steps 10 --no-cpu-baseline > $O/bench_eigen.json 2> $O/bench_eigen.err; echo "bench eigen rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -30 $O/pytest_gpu.log; tail -3 $O/bench_eigen.err; cut -c1-2000 $O/bench_eigen.json
