#!/bin/bash
# 8-GPU check of the bench contract (independent chains, no data-path collective): SBM and eigenmodel workloads.
mkdir -p gpurun_out/val8; O=gpurun_out/val8
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 10 --warmup 3 > $O/bench_sbm_8gpu.json 2> $O/bench_sbm_8gpu.err; echo "sbm8 rc=$?" > $O/rc.txt
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 8 --workload eigen --steps 5 --warmup 3 > $O/bench_eig_8gpu.json 2> $O/bench_eig_8gpu.err; echo "eig8 rc=$?" >> $O/rc.txt
cat $O/rc.txt
