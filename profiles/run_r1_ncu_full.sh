#!/bin/bash
# ncu --set full captures of the dominant kernels of the SBM and eigenmodel sweeps (steady state: launches skipped
# past the burn-in sweeps), summarised with profiles/summarize.py.  Outputs in gpurun_out/full/.
mkdir -p gpurun_out/full; O=gpurun_out/full
ncu --set full --clock-control none --import-source on -k regex:"k_strip_stats|k_sbm_seq2|k_lp_from_stats" -s 140 -c 6 -o $O/sbm_full -f \
    python bench.py --steps 2 --warmup 6 --no-cpu-baseline > $O/ncu_sbm.log 2>&1; echo "ncu sbm rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"k_eig_pg_tile|k_eig_gfull|k_eig_seq4|k_eig_push" -s 520 -c 8 -o $O/eig_full -f \
    python bench.py --workload eigen --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_eig.log 2>&1; echo "ncu eig rc=$?"
python profiles/summarize.py $O/sbm_full.ncu-rep $O/sbm_full.csv; python profiles/summarize.py $O/eig_full.ncu-rep $O/eig_full.csv
cat $O/sbm_full.csv | cut -c1-200; cat $O/eig_full.csv | cut -c1-200
