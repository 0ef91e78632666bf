#!/bin/bash
O=gpurun_out
mkdir -p $O
: > $O/r2_sweep7.txt
for kp in "25 4 128" "26 4 128" "27 4 128" "28 4 128" "29 4 128" "28 8 64" "25 5 128" "24 4 128" "32 4 128"; do set -- $kp; python tools/sweep.py --k $1 --p $2 --frag-mib $3 --variants 0,6 --reps 10 >> $O/r2_sweep7.txt 2>&1; done
grep RS $O/r2_sweep7.txt
timeout 900 python -m pytest tests -m gpu -x -q -k "bank_replicated or peeled or fuzz or batch" > $O/r2_pytest7.log 2>&1; tail -3 $O/r2_pytest7.log
