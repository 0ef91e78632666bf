#!/bin/bash
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "unaligned or guard or peeled" > $O/r2_pytest23.log 2>&1; tail -2 $O/r2_pytest23.log
: > $O/r2_sweep23.txt
python tools/sweep.py --k 10 --p 4 --frag-mib 400 --extra-bytes 3 --variants 0 --reps 10 >> $O/r2_sweep23.txt 2>&1
for kp in "4 2 512" "16 4 256" "12 4 256" "20 4 128" "10 5 256" "32 8 128" "28 4 128" "64 4 64"; do set -- $kp; python tools/sweep.py --k $1 --p $2 --frag-mib $3 --extra-bytes 5 --variants 0 --reps 5 >> $O/r2_sweep23.txt 2>&1; done
grep RS $O/r2_sweep23.txt
