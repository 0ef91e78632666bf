#!/bin/bash
# round-2 call 6 (eight GPUs): what the host links give 4 and 8 concurrent GPUs, bench --gpus 8 with the multi-GPU checks,
# the file front end (C1, C5) with 1 / 2 / 8 workers
O=gpurun_out
mkdir -p $O
nvidia-smi -L > $O/r2_8gpu_smi.txt; nproc >> $O/r2_8gpu_smi.txt; free -g | head -2 >> $O/r2_8gpu_smi.txt
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
CUDA_VISIBLE_DEVICES=0,1,2,3 $TR --nproc-per-node 4 --master-port 29504 tools/pcie_probe.py --json $O/pcie_4.json 2>/dev/null
$TR --nproc-per-node 8 --master-port 29508 tools/pcie_probe.py --json $O/pcie_8.json 2>/dev/null
$TR --nproc-per-node 8 --master-port 29518 bench.py --gpus 8 --steps 20 --warmup 3 > $O/r2_bench_n8.json 2> $O/r2_bench_n8.err; echo "bench n8 rc=$?"
CUDA_VISIBLE_DEVICES=0,1,2,3 $TR --nproc-per-node 4 --master-port 29514 bench.py --gpus 4 --steps 20 --warmup 3 --no-legs > $O/r2_bench_n4.json 2> $O/r2_bench_n4.err; echo "bench n4 rc=$?"
python tests/tools/run_files.py c1 > $O/r2_files_c1.json 2>&1; cat $O/r2_files_c1.json
RS_TIMING=1 python tests/tools/run_files.py big --k 32 --p 8 --gib 8 > $O/r2_files_c5_default.json 2>&1; tail -1 $O/r2_files_c5_default.json
RS_TIMING=1 RS_GPUS=2 python tests/tools/run_files.py big --k 32 --p 8 --gib 8 > $O/r2_files_c5_2gpu.json 2>&1; tail -1 $O/r2_files_c5_2gpu.json
RS_TIMING=1 RS_GPUS=8 python tests/tools/run_files.py big --k 32 --p 8 --gib 8 > $O/r2_files_c5_8gpu.json 2>&1; tail -1 $O/r2_files_c5_8gpu.json
echo "--- all 8 visible" > $O/r2_startup_8gpu.txt; python tools/startup_probe.py >> $O/r2_startup_8gpu.txt 2>&1
echo "--- CUDA_VISIBLE_DEVICES=0" >> $O/r2_startup_8gpu.txt; CUDA_VISIBLE_DEVICES=0 python tools/startup_probe.py >> $O/r2_startup_8gpu.txt 2>&1
cat $O/r2_startup_8gpu.txt
