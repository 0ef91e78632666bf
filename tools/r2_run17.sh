#!/bin/bash
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "plain_xor or decode_roundtrip or repair or encode_vs_oracle or host_pipeline or cli" > $O/r2_pytest17.log 2>&1; tail -3 $O/r2_pytest17.log
: > $O/r2_sweep17.txt
for kp in "10 1 400" "4 1 512" "16 1 256" "32 1 128" "64 1 64"; do set -- $kp; python tools/sweep.py --k $1 --p $2 --frag-mib $3 --variants 0,5 --reps 10 >> $O/r2_sweep17.txt 2>&1; done
grep RS $O/r2_sweep17.txt
