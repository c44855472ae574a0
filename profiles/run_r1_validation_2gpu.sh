mkdir -p gpurun_out/val2; O=gpurun_out/val2
timeout 300 python -m pytest tests/test_gpu_sharded.py -q -m gpu > $O/pytest_sharded.log 2>&1; echo "pytest sharded rc=$?" > $O/rc.txt
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > $O/bench_sbm_2gpu.json 2> $O/bench_sbm_2gpu.err; echo "sbm2 rc=$?" >> $O/rc.txt
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload eigen --steps 10 --warmup 3 > $O/bench_eig_2gpu.json 2> $O/bench_eig_2gpu.err; echo "eig2 rc=$?" >> $O/rc.txt
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus 2 --steps 1 --warmup 0 > $O/bench_ref_2gpu.json 2> $O/bench_ref_2gpu.err; echo "ref2 rc=$?" >> $O/rc.txt
cat $O/rc.txt
