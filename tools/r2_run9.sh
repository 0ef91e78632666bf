#!/bin/bash
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -k "multipass" > $O/r2_pytest9.log 2>&1; tail -2 $O/r2_pytest9.log
: > $O/r2_sweep9.txt
for kp in "28 4 128" "40 4 64" "64 4 64" "128 4 32"; do set -- $kp; python tools/sweep.py --k $1 --p $2 --frag-mib $3 --variants 9,7 --reps 10 >> $O/r2_sweep9.txt 2>&1; done
grep RS $O/r2_sweep9.txt
