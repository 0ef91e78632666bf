#!/bin/bash
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "unaligned or guard or peeled or plain_xor or repair" > $O/r2_pytest19.log 2>&1; tail -3 $O/r2_pytest19.log
: > $O/r2_sweep19.txt
for kp in "28 4 128" "25 4 128" "64 4 64" "40 5 64" "64 8 32"; do set -- $kp; python tools/sweep.py --k $1 --p $2 --frag-mib $3 --extra-bytes 3 --variants 0,6 --reps 5 >> $O/r2_sweep19.txt 2>&1; done
grep RS $O/r2_sweep19.txt
