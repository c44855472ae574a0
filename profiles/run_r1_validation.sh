#!/bin/bash
# Round-1 validation pass on one B200: GPU parity tests, smoke, the three bench workloads, the reference arm and
# ncu launch lists of the same bench commands.  Outputs land in gpurun_out/val/.
mkdir -p gpurun_out/val; O=gpurun_out/val
python -m pytest tests -q -m gpu -x > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" > $O/rc.txt
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/rc.txt
python bench.py > $O/bench_sbm.json 2> $O/bench_sbm.err; echo "bench sbm rc=$?" >> $O/rc.txt
python bench.py --workload eigen --steps 10 > $O/bench_eigen.json 2> $O/bench_eigen.err; echo "bench eigen rc=$?" >> $O/rc.txt
python bench.py --workload latent --steps 10 > $O/bench_latent.json 2> $O/bench_latent.err; echo "bench latent rc=$?" >> $O/rc.txt
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "bench ref rc=$?" >> $O/rc.txt
ncu --metrics gpu__time_duration.sum,sm__cycles_elapsed.max --clock-control none -c 2000 --csv --log-file $O/sbm_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_sbm.log 2>&1; echo "ncu sbm rc=$?" >> $O/rc.txt
ncu --metrics gpu__time_duration.sum,sm__cycles_elapsed.max --clock-control none -c 3000 --csv --log-file $O/eig_launches.csv \
    python bench.py --workload eigen --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_eig.log 2>&1; echo "ncu eig rc=$?" >> $O/rc.txt
python profiles/time_mf.py > $O/time_mf.log 2>&1; echo "time_mf rc=$?" >> $O/rc.txt
cat $O/rc.txt
