#!/bin/bash
# round 2, call m: default bench line with the AIS record; ncu --set full of k_ais_rowcnt / k_ais_seq (source counters)
mkdir -p gpurun_out/r2m
( time timeout 900 python bench.py ) > gpurun_out/r2m/bench_default.json 2> gpurun_out/r2m/bench_default.err; tail -4 gpurun_out/r2m/bench_default.err; cut -c1-300 gpurun_out/r2m/bench_default.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_ais_rowcnt|k_ais_seq" -s 200 -c 4 -o gpurun_out/r2m/ais_full python profiles/time_ais.py 65536 128 > gpurun_out/r2m/ncu_full.log 2>&1; tail -2 gpurun_out/r2m/ncu_full.log
