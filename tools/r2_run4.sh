#!/bin/bash
# round-2 call 4: peeled-row shapes after the block-size fix, e2e pipeline knobs on one GPU, ncu of the realigning kernel
O=gpurun_out
mkdir -p $O
: > $O/r2_sweep4.txt
for kp in "15 5 128" "9 5 256" "24 5 128" "31 5 64"; do set -- $kp; python tools/sweep.py --k $1 --p $2 --frag-mib $3 --variants 0 --reps 10 >> $O/r2_sweep4.txt 2>&1; done
grep RS $O/r2_sweep4.txt
python tools/pcie_probe.py --json $O/pcie_1.json > $O/r2_pcie1.txt 2>&1; cat $O/r2_pcie1.txt
: > $O/r2_e2e_knobs.txt
for env in "X=1" "RSGPU_PIPE_ROWCOPY=1" "RSGPU_PIPE_SLICE_MIB=2" "RSGPU_PIPE_SLICE_MIB=16" "RSGPU_PIPE_SLICE_MIB=32"; do
  for slots in 2 4 8; do
    echo "== env [$env] slots $slots" >> $O/r2_e2e_knobs.txt
    env $env python bench.py --steps 5 --warmup 3 --no-cpu --no-legs --sustained-s 0 --e2e-slots $slots 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['e2e']['value'], d['config']['self_check']['status'])" >> $O/r2_e2e_knobs.txt
  done
done
cat $O/r2_e2e_knobs.txt
python tools/sweep.py --k 10 --p 4 --frag-mib 400 --extra-bytes 3 --variants 0 --reps 3 > $O/plain_ua.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_gf_matmul_fat_ua -s 3 -c 1 -o $O/r2_ua_rs1014 python tools/sweep.py --k 10 --p 4 --frag-mib 400 --extra-bytes 3 --variants 0 --reps 3 > $O/ncu_ua.log 2>&1
python tools/ncu_summary.py $O/r2_ua_rs1014.ncu-rep $O/r2_ncu_full_ua_rs1014.csv "round 2: realigning kernel, RS(10,14), L = 400 MiB + 3, pitch = L" > /dev/null
python tools/sweep.py --k 15 --p 5 --frag-mib 128 --variants 0 --reps 3 > $O/plain_xr.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_gf_matmul_fat -s 3 -c 1 -o $O/r2_xr_rs1520 python tools/sweep.py --k 15 --p 5 --frag-mib 128 --variants 0 --reps 3 > $O/ncu_xr.log 2>&1
python tools/ncu_summary.py $O/r2_xr_rs1520.ncu-rep $O/r2_ncu_full_xr_rs1520.csv "round 2: peeled XOR row, RS(15,20) on the k = 16 shape" > /dev/null
