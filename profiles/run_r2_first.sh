#!/bin/bash
# Round-2 first pass on one B200: GPU parity tests at HEAD, then ncu --set full (source-level) of the serial eigen kernel
# and the Polya-gamma kernel to guide the rewrite.  Outputs in gpurun_out/r2a/.
mkdir -p gpurun_out/r2a; O=gpurun_out/r2a
timeout 900 python -m pytest tests -q -m gpu -x > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" > $O/rc.txt
timeout 300 python bench.py --workload eigen --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_eigen.json 2> $O/bench_eigen.err; echo "bench eigen rc=$?" >> $O/rc.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_eig_seq4|k_eig_pg_tile|k_eig_push" -s 300 -c 5 -o $O/eig_full -f \
    python bench.py --workload eigen --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_eig.log 2>&1; echo "ncu eig rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -3 $O/pytest_gpu.log; cat $O/bench_eigen.json | cut -c1-600
