#!/usr/bin/env python3
"""Kernel tuning sweep (developer tool, GPU box): times rsgpu_encode for variant x threads x grid cap.

    python tools/sweep.py [--k 10 --p 4 --frag-mib 400] [--variants 0,1,2] [--threads 128,256] [--blocks 0]
Prints one line per combination: kernel ms (CUDA events, mean of reps), input GB/s, traffic GB/s,
fraction of the measured HBM peak.
"""
import argparse
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--p", type=int, default=4)
    ap.add_argument("--frag-mib", type=float, default=400)
    ap.add_argument("--variants", default="0,5,6")
    ap.add_argument("--threads", default="0")
    ap.add_argument("--blocks", default="0")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--extra-bytes", type=int, default=0, help="add this many bytes to the fragment length; pitch = length (unaligned rows)")
    args = ap.parse_args()
    import torch
    rs = importlib.import_module("gpu-rscode_b200")
    k, p = args.k, args.p
    L = int(args.frag_mib * (1 << 20)) // 256 * 256 + args.extra_bytes
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        peak = 6650.0
    data = torch.randint(0, 256, (k, L), dtype=torch.uint8, device="cuda")
    out = torch.empty((p, L), dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    # reference point: plain device copy of the same number of bytes (read n*L/2, write n*L/2)
    nb = (k + p) * L // 2
    a = torch.empty(nb, dtype=torch.uint8, device="cuda")
    b = torch.empty(nb, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        b.copy_(a)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.reps):
        b.copy_(a)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.reps
    print(f"copy_ of {nb/1e9:.2f} GB: {ms:.3f} ms  {2*nb/ms/1e6:.0f} GB/s (read+write)  [peak file: {peak}]", flush=True)
    del a, b
    with rs.RSContext(k, p) as ctx:
        for v in [int(x) for x in args.variants.split(",")]:
            for th in [int(x) for x in args.threads.split(",")]:
                for mb in [int(x) for x in args.blocks.split(",")]:
                    ctx.set_tuning(variant=v, threads=th, max_blocks=mb)
                    for _ in range(3):
                        ctx.encode(data, L, out, L, L, st)
                    e0.record()
                    for _ in range(args.reps):
                        ctx.encode(data, L, out, L, L, st)
                    e1.record()
                    torch.cuda.synchronize()
                    ms = e0.elapsed_time(e1) / args.reps
                    tr = (k + p) * L / ms / 1e6
                    print(f"RS({k},{k+p}) L={L} variant={v} threads={th} max_blocks={mb}: {ms:.3f} ms  "
                          f"input {k*L/ms/1e6:.0f} GB/s  traffic {tr:.0f} GB/s  frac {tr/peak:.3f}", flush=True)


if __name__ == "__main__":
    main()
