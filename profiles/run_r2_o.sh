#!/bin/bash
# round 2, call o (2 GPUs): row-sharded strong-scaling state before the device-side label exchange
mkdir -p gpurun_out/r2o
for wl in latent eigen sbm; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --workload $wl --shard rows --no-cpu-baseline > gpurun_out/r2o/rows_${wl}_2gpu.json 2> gpurun_out/r2o/rows_${wl}_2gpu.err
  echo "$wl rc=$?"; cut -c1-260 gpurun_out/r2o/rows_${wl}_2gpu.json; tail -2 gpurun_out/r2o/rows_${wl}_2gpu.err
done
