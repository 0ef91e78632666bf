#!/bin/bash
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node 2 --master-port 29521 bench.py --gpus 2 --steps 20 --warmup 3 > $O/r2_bench_n2_final.json 2> $O/r2_bench_n2_final.err; echo "bench n2 rc=$?"
$TR --nproc-per-node 2 --master-port 29522 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > $O/r2_bench_ref_n2.json 2>/dev/null; echo "ref n2 rc=$?"; cat $O/r2_bench_ref_n2.json | cut -c1-200
timeout 900 python -m pytest tests -m gpu -x -q -k "peer or cli or sharding or guard" > $O/r2_pytest_2gpu_final.log 2>&1; tail -2 $O/r2_pytest_2gpu_final.log
