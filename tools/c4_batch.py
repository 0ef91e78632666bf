#!/usr/bin/env python
"""BASELINE configs[3] at its full size: a batch of 64 RS(16,20) stripes of
16 x 64 MiB (64 GiB of input + 16 GiB of parity), all device resident, GPU g owning the stripes
s = g (mod G) -- strong scaling, no exchange between GPUs.  One step = one pass over the whole batch
(one rsgpu_encode_batch launch per GPU; --per-stripe: one launch per stripe).  G = 1 keeps all 80 GiB on the one GPU (fits 180 GB; no half-batches).

    python tools/c4_batch.py                                   # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29517 tools/c4_batch.py                  # 8 GPUs, 8 stripes each

Prints one JSON line on rank 0.  Every parity-0 fragment is checked against the XOR of its stripe's natives
(generator row 0 is all ones) and one stripe per rank is decoded back after erasing four natives.
"""
import argparse
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

MIB = 1 << 20


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stripes", type=int, default=64)
    ap.add_argument("--frag-mib", type=int, default=64)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--per-stripe", action="store_true", help="one rsgpu_encode launch per stripe instead of one batch call")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rs = importlib.import_module("gpu-rscode_b200")
    rs.lib()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    k, p, L = 16, 4, args.frag_mib * MIB
    mine = [s for s in range(args.stripes) if s % world == rank]
    stream = torch.cuda.current_stream().cuda_stream
    g = torch.Generator(device=dev).manual_seed(0xC4 + rank)
    data = torch.empty((len(mine), k, L), dtype=torch.uint8, device=dev)
    for i in range(len(mine)):                       # stripe by stripe: no 64 GiB temporaries
        data[i].random_(0, 256, generator=g)
    parity = torch.empty((len(mine), p, L), dtype=torch.uint8, device=dev)
    ctx = rs.RSContext(k, p, device=local)

    def step():
        if args.per_stripe:
            for i in range(len(mine)):
                ctx.encode(data[i].data_ptr(), L, parity[i].data_ptr(), L, L, stream)
        elif mine:      # the rank's stripes in ONE launch (rsgpu_encode_batch, stripe-major layout)
            ctx.encode_batch(data.data_ptr(), L, k * L, parity.data_ptr(), L, p * L, L, len(mine), stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        step()
    barrier()
    n0 = ctx.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    launches = ctx.launch_count() - n0
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    ok = True
    for i in range(len(mine)):
        x = data[i, 0].clone()
        for j in range(1, k):
            x ^= data[i, j]
        ok = ok and bool(torch.equal(x, parity[i, 0]))
    if mine:                                        # erase natives 0..3 of the first stripe, decode them back
        order = [16, 17, 18, 19] + list(range(4, 16))
        ctx.set_survivors(order)
        surv = torch.cat([parity[0], data[0, 4:]]).contiguous()
        back = torch.empty((4, L), dtype=torch.uint8, device=dev)
        ctx.decode_erased(surv.data_ptr(), L, back.data_ptr(), L, L, stream)
        torch.cuda.synchronize()
        ok = ok and bool(torch.equal(back, data[0, :4]))
    okt = torch.tensor([1 if ok else 0], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
    if rank == 0:
        ms_step = float(ms.item()) / args.steps
        peak = 6448.7
        try:
            peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        except Exception:
            pass
        in_bytes = args.stripes * k * L
        per_gpu_traffic = len(mine) * (k + p) * L / (ms_step / 1e3) / 1e9
        print(json.dumps({
            "metric": "encode GB/s (input bytes) RS(16,20), %d-stripe batch" % args.stripes, "value": in_bytes / (ms_step / 1e3) / 1e9,
            "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms_step,
            "scaling": "strong", "dtype": "u8", "data": "synthetic",
            "config": {"workload": "BASELINE configs[3]: %d stripes x 16 x %d MiB, stripes s = g (mod G), all device resident"
                                   % (args.stripes, args.frag_mib), "input_bytes": in_bytes,
                       "resident_bytes_per_gpu": len(mine) * (k + p) * L, "self_check": "ok" if int(okt.item()) else "MISMATCH"},
            "roofline": {"bound": "hbm", "achieved": per_gpu_traffic, "peak": peak, "unit": "GB/s", "frac": per_gpu_traffic / peak,
                         "note": "rank 0's stripes, whole step; " + ("one launch per stripe" if args.per_stripe else "one batch launch")},
            "gpu_launches": launches * world}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
