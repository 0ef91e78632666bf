#!/bin/bash
# round-2 call 5 (two GPUs): bench --gpus 2 with the peer / two-worker checks, PCIe probe 1 vs 2 ranks, start-up cost per visible GPU,
# the 2-GPU tests, and the tensor-core probe
O=gpurun_out
mkdir -p $O
nvidia-smi -L > $O/r2_2gpu_smi.txt
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node 2 --master-port 29501 bench.py --gpus 2 --steps 20 --warmup 3 > $O/r2_bench_n2.json 2> $O/r2_bench_n2.err; echo "bench n2 rc=$?"
python tools/pcie_probe.py --json $O/pcie_1of2.json
$TR --nproc-per-node 2 --master-port 29502 tools/pcie_probe.py --json $O/pcie_2.json 2>/dev/null
echo "--- both GPUs visible" > $O/r2_startup_2gpu.txt; python tools/startup_probe.py >> $O/r2_startup_2gpu.txt 2>&1
echo "--- CUDA_VISIBLE_DEVICES=0" >> $O/r2_startup_2gpu.txt; CUDA_VISIBLE_DEVICES=0 python tools/startup_probe.py >> $O/r2_startup_2gpu.txt 2>&1
cat $O/r2_startup_2gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q -k "peer or cli or sharding" > $O/r2_pytest_2gpu.log 2>&1; tail -3 $O/r2_pytest_2gpu.log
tools/tc_bitslice 64 10 > $O/r2_tc_bitslice.txt 2>&1; cat $O/r2_tc_bitslice.txt
python tools/sweep.py --k 32 --p 8 --frag-mib 64 --variants 0 --reps 10 | grep RS >> $O/r2_tc_bitslice.txt
