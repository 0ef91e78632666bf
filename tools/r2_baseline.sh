#!/bin/bash
# round-2 baseline of the round-1 build: profiles the judge asked for + today's figures of what is about to change
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/r2b_smi.txt
python tools/startup_probe.py > $O/r2b_startup.txt 2>&1
python tools/startup_probe.py >> $O/r2b_startup.txt 2>&1
# unaligned rows today (byte kernel): RS(10,14), L = 400 MiB + 3, pitch = L
python tools/sweep.py --k 10 --p 4 --frag-mib 400 --extra-bytes 3 --variants 0 --reps 3 > $O/r2b_unaligned.txt 2>&1
# the codes that miss the target
for kp in "32 8 128" "15 5 128" "28 4 128" "64 4 64" "10 5 256" "16 8 128"; do set -- $kp; python tools/sweep.py --k $1 --p $2 --frag-mib $3 --variants 0 >> $O/r2b_sweep.txt 2>&1; done
# full captures
python tools/sweep.py --k 32 --p 8 --frag-mib 128 --variants 0 --reps 3 > $O/plain_3240.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_gf_matmul -s 3 -c 1 -o $O/r2b_rs3240 python tools/sweep.py --k 32 --p 8 --frag-mib 128 --variants 0 --reps 3 > $O/ncu_3240.log 2>&1
python tools/sweep.py --k 15 --p 5 --frag-mib 128 --variants 0 --reps 3 > $O/plain_1520.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_gf_matmul -s 3 -c 1 -o $O/r2b_rs1520 python tools/sweep.py --k 15 --p 5 --frag-mib 128 --variants 0 --reps 3 > $O/ncu_1520.log 2>&1
python tools/sweep.py --k 16 --p 4 --frag-mib 256 --variants 7 --reps 3 > $O/plain_tma.log 2>&1 && \
ncu --set full --clock-control none -k regex:k_gf_matmul_tma -s 3 -c 1 -o $O/r2b_tma1620 python tools/sweep.py --k 16 --p 4 --frag-mib 256 --variants 7 --reps 3 > $O/ncu_tma.log 2>&1
python tools/sweep.py --k 28 --p 4 --frag-mib 128 --variants 0 --reps 3 > $O/plain_2832.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_gf_matmul -s 3 -c 1 -o $O/r2b_rs2832 python tools/sweep.py --k 28 --p 4 --frag-mib 128 --variants 0 --reps 3 > $O/ncu_2832.log 2>&1
python bench.py --workload c3 --steps 3 --warmup 3 --no-e2e --no-cpu --sustained-s 0 > $O/plain_c3.log 2>&1 && \
ncu --set full --clock-control none -k regex:k_decode_setup -s 2 -c 1 -o $O/r2b_decode_setup python bench.py --workload c3 --steps 3 --warmup 3 --no-e2e --no-cpu --sustained-s 0 > $O/ncu_c3.log 2>&1
ls -la $O/*.ncu-rep
for n in rs3240 rs1520 tma1620 rs2832 decode_setup; do
  [ -f $O/r2b_$n.ncu-rep ] && python tools/ncu_summary.py $O/r2b_$n.ncu-rep $O/r2b_ncu_full_$n.csv "round 2 baseline (round-1 build): $n" > /dev/null
done
du -sh $O
