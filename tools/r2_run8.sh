#!/bin/bash
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -k "multipass or variants" > $O/r2_pytest8.log 2>&1; tail -3 $O/r2_pytest8.log
: > $O/r2_sweep8.txt
for kp in "25 4 128" "28 4 128" "32 4 128" "40 4 64" "64 4 64" "128 4 32" "128 2 32" "32 8 128" "64 8 32" "24 4 128" "20 4 128"; do set -- $kp; python tools/sweep.py --k $1 --p $2 --frag-mib $3 --variants 9,5,6,7 --reps 10 >> $O/r2_sweep8.txt 2>&1; done
grep RS $O/r2_sweep8.txt
python tools/sweep.py --k 64 --p 4 --frag-mib 64 --variants 9 --reps 3 > $O/plain_mp.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_gf_matmul_fat_mp -s 3 -c 1 -o $O/r2_mp_rs6468 python tools/sweep.py --k 64 --p 4 --frag-mib 64 --variants 9 --reps 3 > $O/ncu_mp.log 2>&1
python tools/ncu_summary.py $O/r2_mp_rs6468.ncu-rep $O/r2_ncu_full_mp_rs6468.csv "round 2: multi-pass byte-table kernel, RS(64,68), 64 MiB fragments" | grep -v "per_second\|occupancy_limit"
