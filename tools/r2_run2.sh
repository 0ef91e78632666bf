#!/bin/bash
# round-2 call 2: full GPU test-suite, the new bench line, unaligned-row sweep
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r2_pytest_gpu.log
tail -5 $O/r2_pytest_gpu.log
python bench.py --steps 50 --warmup 5 > $O/r2_bench_c2.json 2> $O/r2_bench_c2.err; echo "bench rc=$?"
python tools/sweep.py --k 10 --p 4 --frag-mib 400 --extra-bytes 3 --variants 0,6 --reps 5 > $O/r2_unaligned.txt 2>&1
python tools/sweep.py --k 10 --p 4 --frag-mib 400 --variants 0 --reps 10 >> $O/r2_unaligned.txt 2>&1
for kp in "4 2 512" "16 4 256" "32 8 128" "15 5 128" "12 4 256" "20 4 128"; do set -- $kp; python tools/sweep.py --k $1 --p $2 --frag-mib $3 --extra-bytes 5 --variants 0 --reps 5 >> $O/r2_unaligned.txt 2>&1; done
cat $O/r2_unaligned.txt | grep RS
python __graft_entry__.py smoke > $O/r2_smoke.log 2>&1; tail -1 $O/r2_smoke.log
