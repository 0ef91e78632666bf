#!/bin/bash
O=gpurun_out
mkdir -p $O
CMD="python bench.py --steps 5 --warmup 3 --no-cpu --sustained-s 0 --no-e2e"
$CMD > $O/plain_bench_launches.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/r2_launches_bench_default.csv $CMD > $O/ncu_bench_launches.log 2>&1
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open('gpurun_out/r2_launches_bench_default.csv')) if len(r)>10]
h=rows[0]; kn=h.index('Kernel Name'); mv=h.index('Metric Value'); mu=h.index('Metric Unit')
tot=collections.defaultdict(lambda:[0,0.0])
for r in rows[1:]:
    try: v=float(r[mv].replace(',',''))
    except: continue
    u=r[mu]; v_us = v/1e3 if u in('ns','nsecond') else (v*1e3 if u in ('ms','msecond') else v)
    name=r[kn][:90]; tot[name][0]+=1; tot[name][1]+=v_us
tt=sum(v[1] for v in tot.values())
print("kernel,launches,total_us,share")
for n,(c,t) in sorted(tot.items(), key=lambda x:-x[1][1])[:25]:
    print(f"\"{n}\",{c},{t:.1f},{t/tt:.4f}")
PY
