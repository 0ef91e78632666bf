#!/usr/bin/env python
"""What the host links give: pinned H2D alone, D2H alone, both at once (two streams) -- the ceiling of the e2e
leg of bench.py (rsgpu_encode_host moves k bytes in and p bytes out per k input bytes).

    python tools/pcie_probe.py [--mib 2048]                                  one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        tools/pcie_probe.py --json gpurun_out/pcie_N.json                    N GPUs at the same time, one process each

With N ranks every rank drives its own GPU from its own pinned buffers, all start together after a barrier, and
the figure is the aggregate: total bytes / the slowest rank's time (max over ranks).  `encode_input_ceiling_gbs`
is what that aggregate allows RS(10,14) encode through host buffers: input rate r with r in + 0.4 r out at once.
"""
import argparse
import json
import os
import time

import torch

ap = argparse.ArgumentParser()
ap.add_argument("--mib", type=int, default=2048)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--json", default="")
a = ap.parse_args()
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = a.mib << 20
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(h2d, d2h, reps, out_frac=1.0):
    """aggregate GB/s per direction over all ranks"""
    m = int(n * out_frac) // 256 * 256
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        if h2d:
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2):
                h_out[:m].copy_(d_out[:m], non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([dt], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    return world * reps * n / dt / 1e9


run(True, True, 1)
h2d = run(True, False, a.reps)
d2h = run(False, True, a.reps)
both = run(True, True, a.reps)
enc = run(True, True, a.reps, out_frac=0.4)      # the RS(10,14) encode mix: 1.0 in, 0.4 out
if rank == 0:
    res = {"n_gpus": world, "buffer_mib": a.mib, "h2d_gbs": h2d, "d2h_gbs": d2h, "h2d_d2h_concurrent_gbs_per_direction": both,
           "encode_input_ceiling_gbs": enc,
           "source": "tools/pcie_probe.py, %d concurrent process(es), pinned %d MiB buffers, aggregate = bytes / slowest rank" % (world, a.mib)}
    print(f"{world} GPU(s): H2D alone {h2d:.1f} GB/s | D2H alone {d2h:.1f} GB/s | both at once {both:.1f} GB/s each direction | "
          f"1.0 in + 0.4 out (RS(10,14) encode mix) {enc:.1f} GB/s of input")
    if a.json:
        json.dump(res, open(a.json, "w"), indent=1)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
