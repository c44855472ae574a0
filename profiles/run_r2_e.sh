#!/bin/bash
mkdir -p gpurun_out/r2e; O=gpurun_out/r2e
timeout 400 python -m pytest tests/test_gpu_eigen.py tests/test_gpu_probit.py tests/test_gpu_config_sizes.py tests/test_gpu_mf_eigen.py -q -m gpu -x > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" > $O/rc.txt
timeout 300 python bench.py --workload eigen --steps 10 --no-cpu-baseline > $O/bench_eigen.json 2> $O/bench_eigen.err; echo "bench eigen rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -3 $O/pytest_gpu.log; python -c "
import json; d=json.load(open('$O/bench_eigen.json')); r=d['roofline']; print(d['value'], d['ms_per_step'], d['e2e']['value'], r['other_kernels_ms_per_sweep'], r['kernel_ms_per_sweep'], r['achieved'], r['frac'])"
