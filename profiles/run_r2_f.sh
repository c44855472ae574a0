#!/bin/bash
mkdir -p gpurun_out/r2f; O=gpurun_out/r2f
python bench.py --workload eigen --steps 1 --warmup 3 --no-cpu-baseline > $O/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_eig_seq5 -s 200 -c 2 -o $O/seq5 -f \
    python bench.py --workload eigen --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu.log 2>&1; echo "ncu rc=$?"
ls -la $O; tail -3 $O/ncu.log
