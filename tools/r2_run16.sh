#!/bin/bash
O=gpurun_out
python tools/sweep.py --k 24 --p 4 --frag-mib 128 --variants 8,5 --reps 3 > $O/plain_cluster.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_gf_matmul_fat_cluster -s 3 -c 1 -o $O/r2_cluster_rs2428 python tools/sweep.py --k 24 --p 4 --frag-mib 128 --variants 8 --reps 3 > $O/ncu_cluster.log 2>&1
python tools/ncu_summary.py $O/r2_cluster_rs2428.ncu-rep $O/r2_ncu_full_cluster_rs2428.csv "round 2: 2-CTA cluster kernel (variant 8), RS(24,28), 128 MiB fragments" | grep -v "per_second\|occupancy_limit"
cat $O/plain_cluster.log | grep RS
ncu -i $O/r2_cluster_rs2428.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); h=rows[0]; d=rows[2]
for i,n in enumerate(h):
    if 'issue_stalled' in n and 'per_issue_active' in n and not n.endswith('_not_issued.ratio'): print(n.split('stalled_')[1], d[i])
"
