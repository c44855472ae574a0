#!/bin/bash
# ncu evidence for the eigenmodel sweep at HEAD: launch list of one sweep + --set full captures of the hot kernels
mkdir -p gpurun_out/r2g; O=gpurun_out/r2g
python bench.py --workload eigen --steps 1 --warmup 3 --no-cpu-baseline > $O/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 1200 -c 420 --csv --log-file $O/eig_launches.csv \
    python bench.py --workload eigen --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu1.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_eig_gfull_tc|k_eig_pg_tile" -s 2 -c 4 -o $O/eig_tc_full -f \
    python bench.py --workload eigen --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu2.log 2>&1; echo "ncu full rc=$?"
ls -la $O
