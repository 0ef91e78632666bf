#!/usr/bin/env python
"""Same-box GPU baseline (optional): the reference's own encode_chunk / decode_chunk
(src/matrix.cu, unmodified, recompiled for sm_100a -> oracle/_ref/libgpurs_ref.so) timed beside this
engine's kernels on the same device buffers.  The reference times itself with CUDA events inside
encode_chunk (matrix.cu:793-823); that number is printed, and an outer CUDA-event bracket around the call
is printed next to it.  Its sizes are `int`, so k * L stays below 2 GiB (L = 200 MiB for RS(10,14)).

    python tests/tools/ref_gpu_bench.py [--reps 10]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402
from gpu_lib import rs  # noqa: E402
from oracle_lib import ReferenceGPU  # noqa: E402


def median(xs):
    xs = sorted(xs)
    return xs[len(xs) // 2]


def timed(fn, reps):
    out = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        out.append(a.elapsed_time(b))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=10)
    args = ap.parse_args()
    peak = 6448.7
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    ref = ReferenceGPU()
    stream = torch.cuda.current_stream().cuda_stream
    # (k, p, L): RS(4,6), RS(10,14) at the largest L the reference's int indexing allows, RS(16,20), RS(32,40)
    shapes = [(4, 2, 256 << 20), (10, 4, 200 << 20), (16, 4, 120 << 20), (32, 8, 60 << 20)]
    for k, p, L in shapes:
        data = torch.randint(0, 256, (k, L), dtype=torch.uint8, device="cuda")
        ours = torch.empty((p, L), dtype=torch.uint8, device="cuda")
        theirs = torch.empty((p, L), dtype=torch.uint8, device="cuda")
        with rs.RSContext(k, p) as ctx:
            coeff = torch.from_numpy(ctx.total_matrix()[k:].copy()).cuda()
            for _ in range(3):
                ctx.encode(data.data_ptr(), L, ours.data_ptr(), L, L, stream)
                ref.encode_chunk(data.data_ptr(), coeff.data_ptr(), theirs.data_ptr(), k, p, L)
            torch.cuda.synchronize()
            t_ours = median(timed(lambda: ctx.encode(data.data_ptr(), L, ours.data_ptr(), L, L, stream), args.reps))
            inner = []
            t_ref_outer = median(timed(lambda: inner.append(
                ref.encode_chunk(data.data_ptr(), coeff.data_ptr(), theirs.data_ptr(), k, p, L)), args.reps))
            t_ref = median(inner)
        same = bool(torch.equal(ours, theirs))
        gb_in = k * L / 1e9
        gb_all = (k + p) * L / 1e9
        print(f"RS({k},{k + p}) L={L >> 20} MiB  same_bytes={same}  "
              f"reference kernel {t_ref:.3f} ms (outer {t_ref_outer:.3f}) = {gb_in / t_ref * 1e3:.1f} GB/s input, "
              f"{gb_all / t_ref * 1e3 / peak:.3f} of HBM peak | this engine {t_ours:.3f} ms = "
              f"{gb_in / t_ours * 1e3:.1f} GB/s input, {gb_all / t_ours * 1e3 / peak:.3f} | x{t_ref / t_ours:.1f}",
              flush=True)
        del data, ours, theirs


if __name__ == "__main__":
    main()
