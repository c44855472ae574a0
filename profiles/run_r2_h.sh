#!/bin/bash
mkdir -p gpurun_out/r2h; O=gpurun_out/r2h
python bench.py --workload eigen --steps 1 --warmup 3 --no-cpu-baseline > $O/plain.log 2>&1 || exit 1
for c in 0 9 16 23; do
  if [ $c = 0 ]; then unset GK_TC_CHUNKS; else export GK_TC_CHUNKS=$c; fi
  timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_eig_gfull_tc -s 2 -c 3 --csv --log-file $O/tc_$c.csv \
    python bench.py --workload eigen --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_$c.log 2>&1
  echo "chunks=$c: $(grep k_eig_gfull_tc $O/tc_$c.csv | awk -F'","' '{print $(NF)}' | tr -d '"' | tr '\n' ' ')"
done
