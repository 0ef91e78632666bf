#!/bin/bash
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q -k "peeled or unaligned or bank_replicated or variants or rows or batch or host_pipeline or no_alloc" > $O/r2_pytest3.log 2>&1; echo "pytest rc=$?" >> $O/r2_pytest3.log
tail -4 $O/r2_pytest3.log
: > $O/r2_sweep3.txt
python tools/sweep.py --k 10 --p 4 --frag-mib 400 --extra-bytes 3 --variants 0 --reps 5 >> $O/r2_sweep3.txt 2>&1
for kp in "4 2 512" "16 4 256" "32 8 128" "12 4 256" "20 4 128" "10 5 256"; do set -- $kp; python tools/sweep.py --k $1 --p $2 --frag-mib $3 --extra-bytes 5 --variants 0 --reps 5 >> $O/r2_sweep3.txt 2>&1; done
for kp in "15 5 128" "10 5 256" "6 5 256" "20 5 128" "10 13 128"; do set -- $kp; python tools/sweep.py --k $1 --p $2 --frag-mib $3 --variants 0 --reps 10 >> $O/r2_sweep3.txt 2>&1; done
grep RS $O/r2_sweep3.txt
