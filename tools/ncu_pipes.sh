#!/bin/bash
# usage: tools/ncu_pipes.sh <variant> [extra sweep args]  -- pipe utilisation of the multiply kernel (GPU box)
v=$1; shift
M=gpu__time_duration.sum,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fmalite.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum
python tools/sweep.py --variants $v --reps 3 "$@" > gpurun_out/plain_v$v.log 2>&1 && \
ncu --metrics $M --clock-control none -k regex:k_gf_matmul -s 3 -c 1 --csv --log-file gpurun_out/pipes_v$v.csv python tools/sweep.py --variants $v --reps 3 "$@" > gpurun_out/ncu_v$v.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open('gpurun_out/pipes_v$v.csv')) if len(r)>10]
h=rows[0]; 
for r in rows[1:]:
    print('v$v', r[h.index('Metric Name')], r[h.index('Metric Value')])
PY
