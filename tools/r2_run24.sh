#!/bin/bash
O=gpurun_out
python tools/sweep.py --k 32 --p 1 --frag-mib 128 --variants 0 --reps 3 > $O/plain_xor.log 2>&1 && \
ncu --set full --clock-control none -k regex:k_xor_rows -s 3 -c 1 -o $O/r2_xor_rs3233 python tools/sweep.py --k 32 --p 1 --frag-mib 128 --variants 0 --reps 3 > $O/ncu_xor.log 2>&1
python tools/ncu_summary.py $O/r2_xor_rs3233.ncu-rep $O/r2_ncu_full_xor_rs3233.csv "round 2: XOR kernel (single all-ones row), RS(32,33), 128 MiB fragments" | grep -E "time_duration|dram__bytes|dram_throughput|issue_active|registers"
python tools/sweep.py --k 28 --p 4 --frag-mib 128 --extra-bytes 3 --variants 0 --reps 3 > $O/plain_uan.log 2>&1 && \
ncu --set full --clock-control none -k regex:k_gf_matmul_ua -s 3 -c 1 -o $O/r2_uan_rs2832 python tools/sweep.py --k 28 --p 4 --frag-mib 128 --extra-bytes 3 --variants 0 --reps 3 > $O/ncu_uan.log 2>&1
python tools/ncu_summary.py $O/r2_uan_rs2832.ncu-rep $O/r2_ncu_full_ua_nibble_rs2832.csv "round 2: realigning split-nibble kernel, RS(28,32), L = 128 MiB + 3, pitch = L" | grep -E "time_duration|dram__bytes|pipe_alu|pipe_lsu|wavefronts|issue_active|registers"
