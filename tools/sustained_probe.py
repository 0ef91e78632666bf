#!/usr/bin/env python3
"""Sustained (power-capped) rate of rsgpu_encode against the grid cap: does running fewer SMs at a higher clock pay?
    python tools/sustained_probe.py [--k 10 --p 4 --frag-mib 400] [--blocks 0,140,132,120,104] [--seconds 2.5]
Per setting: 1.5 s of back-to-back launches to get onto the board's power cap, then `seconds` timed with CUDA events;
nvidia-smi sampled alongside (SM clock, power)."""
import argparse
import importlib
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ap = argparse.ArgumentParser()
ap.add_argument("--k", type=int, default=10)
ap.add_argument("--p", type=int, default=4)
ap.add_argument("--frag-mib", type=float, default=400)
ap.add_argument("--blocks", default="0,140,132,120,104")
ap.add_argument("--threads", default="0")
ap.add_argument("--seconds", type=float, default=2.5)
a = ap.parse_args()
import torch  # noqa: E402
rs = importlib.import_module("gpu-rscode_b200")
k, p = a.k, a.p
L = int(a.frag_mib * (1 << 20)) // 256 * 256
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    peak = 6650.0
data = torch.randint(0, 256, (k, L), dtype=torch.uint8, device="cuda")
out = torch.empty((p, L), dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with rs.RSContext(k, p) as ctx:
    for th in [int(x) for x in a.threads.split(",")]:
        for mb in [int(x) for x in a.blocks.split(",")]:
            ctx.set_tuning(variant=0, threads=th, max_blocks=mb)
            for _ in range(5):
                ctx.encode(data, L, out, L, L, st)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(20):
                ctx.encode(data, L, out, L, L, st)
            e1.record()
            torch.cuda.synchronize()
            burst = e0.elapsed_time(e1) / 20
            n_warm = int(1500 / burst)
            for _ in range(n_warm):
                ctx.encode(data, L, out, L, L, st)
            smi = subprocess.Popen(["nvidia-smi", "-i", "0", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits", "-lms", "100"],
                                   stdout=subprocess.PIPE, text=True)
            n = int(a.seconds * 1e3 / burst)
            e0.record()
            for _ in range(n):
                ctx.encode(data, L, out, L, L, st)
            e1.record()
            torch.cuda.synchronize()
            smi.terminate()
            lines = [ln.split(",") for ln in smi.communicate()[0].splitlines() if "," in ln]
            clk = sorted(float(x[0]) for x in lines)
            pw = sorted(float(x[1]) for x in lines)
            ms = e0.elapsed_time(e1) / n
            tr = (k + p) * L / ms / 1e6
            print(f"RS({k},{k+p}) threads={th} max_blocks={mb}: burst {burst:.3f} ms ({(k+p)*L/burst/1e6/peak:.3f})  sustained {ms:.3f} ms  "
                  f"traffic {tr:.0f} GB/s  frac {tr/peak:.3f}  sm clock median {clk[len(clk)//2] if clk else 0:.0f} MHz  power median {pw[len(pw)//2] if pw else 0:.0f} W",
                  flush=True)
            torch.cuda.synchronize()
