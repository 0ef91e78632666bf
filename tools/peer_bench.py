#!/usr/bin/env python3
"""rsgpu_encode_rows with the parity rows on peer GPUs (2+ GPUs): input GB/s and NVLink GB/s.
    python tools/peer_bench.py [--k 10 --p 4 --frag-mib 256]"""
import argparse, importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
rs = importlib.import_module("gpu-rscode_b200")
ap = argparse.ArgumentParser()
ap.add_argument("--k", type=int, default=10); ap.add_argument("--p", type=int, default=4)
ap.add_argument("--frag-mib", type=float, default=256); ap.add_argument("--reps", type=int, default=20)
a = ap.parse_args()
k, p, L = a.k, a.p, int(a.frag_mib * (1 << 20))
n = torch.cuda.device_count()
torch.cuda.set_device(0)
data = torch.randint(0, 256, (k, L), dtype=torch.uint8, device="cuda:0")
st = torch.cuda.current_stream().cuda_stream
with rs.RSContext(k, p, device=0) as ctx:
    for d in range(1, n):
        ctx.enable_peer(d)
    for name, place in [("all rows local", lambda i: 0), ("rows on peers (round robin over GPUs 1..)", lambda i: 1 + i % max(1, n - 1) if n > 1 else 0)]:
        rows = [torch.empty(L, dtype=torch.uint8, device=f"cuda:{place(i)}") for i in range(p)]
        for _ in range(3):
            ctx.encode_rows(data, L, rows, L, st)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.reps):
            ctx.encode_rows(data, L, rows, L, st)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.reps
        remote = sum(1 for i in range(p) if place(i) != 0) * L
        print(f"RS({k},{k+p}) L={L} {name}: {ms:.3f} ms  input {k*L/ms/1e6:.0f} GB/s  peer writes {remote/ms/1e6:.0f} GB/s over {len(set(place(i) for i in range(p)) - {0})} link target(s)")
