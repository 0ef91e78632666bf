#!/bin/bash
# final rehearsal of the driver's scaling run on the 8-GPU box + one file-pipeline experiment (IO threads on 32 vCPUs)
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node 8 --master-port 29531 bench.py --gpus 8 --steps 20 --warmup 3 > $O/r2_bench_n8_final.json 2> $O/r2_bench_n8_final.err; echo "bench n8 rc=$?"
CUDA_VISIBLE_DEVICES=0,1,2,3 $TR --nproc-per-node 4 --master-port 29532 bench.py --gpus 4 --steps 20 --warmup 3 > $O/r2_bench_n4_final.json 2> $O/r2_bench_n4_final.err; echo "bench n4 rc=$?"
RS_TIMING=1 RS_IO_THREADS=32 python tests/tools/run_files.py big --k 32 --p 8 --gib 8 > $O/r2_files_c5_io32.json 2>&1; tail -1 $O/r2_files_c5_io32.json | cut -c1-400
