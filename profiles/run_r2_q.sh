#!/bin/bash
# round 2, call q (2 GPUs): host-timer table of the SBM sweep, one process alone vs two processes side by side
mkdir -p gpurun_out/r2q
nproc; lscpu | grep -E "Model name|^CPU\(s\)|Thread|MHz" | head -6
CUDA_VISIBLE_DEVICES=0 python profiles/prof_sbm_host.py > gpurun_out/r2q/alone.log 2>&1
grep -E "steady|sweep [0-5]" gpurun_out/r2q/alone.log | tail -3
CUDA_VISIBLE_DEVICES=0 OMP_NUM_THREADS=1 python profiles/prof_sbm_host.py > gpurun_out/r2q/alone_omp1.log 2>&1
grep -E "steady" gpurun_out/r2q/alone_omp1.log
CUDA_VISIBLE_DEVICES=0 python profiles/prof_sbm_host.py > gpurun_out/r2q/pair0.log 2>&1 &
CUDA_VISIBLE_DEVICES=1 python profiles/prof_sbm_host.py > gpurun_out/r2q/pair1.log 2>&1 &
wait
grep -E "steady" gpurun_out/r2q/pair0.log gpurun_out/r2q/pair1.log
head -30 gpurun_out/r2q/pair0.log | tail -22
