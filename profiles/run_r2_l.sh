#!/bin/bash
# round 2, call l: AIS bench workload (1 GPU), smoke, ncu launch list + --set full of the two AIS kernels
mkdir -p gpurun_out/r2l
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2l/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/r2l/smoke.log; tail -2 gpurun_out/r2l/smoke.log
timeout 600 python bench.py --workload ais --steps 6 --warmup 3 > gpurun_out/r2l/bench_ais_p128.json 2> gpurun_out/r2l/bench_ais_p128.err; tail -c 1500 gpurun_out/r2l/bench_ais_p128.json; tail -3 gpurun_out/r2l/bench_ais_p128.err
timeout 600 python bench.py --workload ais --steps 6 --warmup 3 --ais-particles 512 --no-cpu-baseline > gpurun_out/r2l/bench_ais_p512.json 2> gpurun_out/r2l/bench_ais_p512.err; cut -c1-400 gpurun_out/r2l/bench_ais_p512.json
timeout 600 python bench.py --workload ais --steps 6 --warmup 3 --ais-particles 8 --no-cpu-baseline > gpurun_out/r2l/bench_ais_p8.json 2> gpurun_out/r2l/bench_ais_p8.err; cut -c1-400 gpurun_out/r2l/bench_ais_p8.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_ais_ -s 40 -c 4 -o gpurun_out/r2l/ais_full python profiles/time_ais.py 65536 128 > gpurun_out/r2l/ncu_full.log 2>&1; tail -2 gpurun_out/r2l/ncu_full.log
