#!/usr/bin/env python
"""Small stripes: one rsgpu_encode call per stripe against ONE rsgpu_encode_batch call for the whole batch,
same device buffers (stripe-major, [stripe][fragment][L]), same total bytes.  CUDA-event times, B200.

    python tools/batch_bench.py [--k 10 --p 4 --total-mib 2048]
"""
import argparse
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


def timed(fn, reps):
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--p", type=int, default=4)
    ap.add_argument("--total-mib", type=int, default=2048, help="input bytes of the whole batch")
    ap.add_argument("--reps", type=int, default=7)
    args = ap.parse_args()
    rs = importlib.import_module("gpu-rscode_b200")
    rs.lib()
    peak = 6448.7
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    k, p = args.k, args.p
    stream = torch.cuda.current_stream().cuda_stream
    total = args.total_mib << 20
    data = torch.randint(0, 256, (total,), dtype=torch.uint8, device="cuda")
    parity_a = torch.empty((total // k * p,), dtype=torch.uint8, device="cuda")
    parity_b = torch.empty_like(parity_a)
    with rs.RSContext(k, p) as ctx:
        for L in (64 << 10, 256 << 10, 1 << 20, 4 << 20, 16 << 20, 64 << 20):
            n = total // (k * L)
            if n < 1:
                continue

            def per_stripe():
                for s in range(n):
                    ctx.encode(data.data_ptr() + s * k * L, L, parity_a.data_ptr() + s * p * L, L, L, stream)

            def batch():
                ctx.encode_batch(data.data_ptr(), L, k * L, parity_b.data_ptr(), L, p * L, L, n, stream)
            per_stripe()
            batch()
            torch.cuda.synchronize()
            same = bool(torch.equal(parity_a[: n * p * L], parity_b[: n * p * L]))
            t_a, t_b = timed(per_stripe, args.reps), timed(batch, args.reps)
            traffic = n * (k + p) * L / 1e9
            print(f"RS({k},{k + p}) fragments of {L >> 10:6d} KiB x {n:5d} stripes  same_bytes={same}  "
                  f"per-stripe launches {t_a:8.3f} ms = {traffic / t_a * 1e3 / peak:.3f} of HBM peak | "
                  f"one batch call {t_b:7.3f} ms = {traffic / t_b * 1e3 / peak:.3f} | x{t_a / t_b:.2f}", flush=True)


if __name__ == "__main__":
    main()
