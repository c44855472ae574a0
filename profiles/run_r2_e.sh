#!/bin/bash
mkdir -p gpurun_out/r2e; O=gpurun_out/r2e
timeout 300 python -m pytest tests/test_gpu_eigen.py tests/test_gpu_probit.py -q -m gpu -x > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" > $O/rc.txt
GK_EIG_PROF=1 timeout 300 python bench.py --workload eigen --steps 3 --no-cpu-baseline > $O/bench_prof.json 2> $O/bench_prof.err; echo "prof rc=$?" >> $O/rc.txt
timeout 300 python bench.py --workload eigen --steps 10 --no-cpu-baseline > $O/bench_eigen.json 2> $O/bench_eigen.err; echo "bench eigen rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -3 $O/pytest_gpu.log; grep "seq5 prof" $O/bench_prof.err | tail -1; python -c "
import json; d=json.load(open('$O/bench_eigen.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['other_kernels_ms_per_sweep'], d['roofline']['kernel_ms_per_sweep'])"
