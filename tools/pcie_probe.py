#!/usr/bin/env python
"""What the host link gives: pinned H2D alone, D2H alone, both at once (two streams) -- the ceiling of the e2e
leg of bench.py (rsgpu_encode_host moves k bytes in and p bytes out per k input bytes).
    python tools/pcie_probe.py [--mib 2048]"""
import argparse
import time

import torch

ap = argparse.ArgumentParser()
ap.add_argument("--mib", type=int, default=2048)
a = ap.parse_args()
n = a.mib << 20
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(h2d, d2h, reps=5):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        if h2d:
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize()
    return reps * n / (time.perf_counter() - t0) / 1e9


run(True, True, 1)
print(f"H2D alone {run(True, False):.1f} GB/s | D2H alone {run(False, True):.1f} GB/s | both at once {run(True, True):.1f} GB/s each direction")
