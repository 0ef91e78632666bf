#!/usr/bin/env python3
"""Small run of every kernel family for compute-sanitizer (GPU box):
    compute-sanitizer --tool memcheck python tests/tools/sanitize_smoke.py
Uses raw cudaMalloc'd buffers through the C ABI (no torch), exact-size allocations so that any
out-of-bounds access of the kernels is visible to the tool, and checks results against the oracle."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
rs = importlib.import_module("gpu-rscode_b200")
from oracle_lib import Oracle  # noqa: E402

orc = Oracle()
rng = np.random.default_rng(0)
cases = [  # k, p, L, variant
    (10, 4, 16 * 3000 + 5, 0), (10, 4, 16 * 3000, 5), (16, 4, 16 * 2049, 5), (4, 2, 16 * 777, 5), (32, 8, 16 * 1500 + 3, 5),
    (11, 3, 16 * 1000 + 1, 0), (13, 5, 4099, 0), (255, 8, 512, 0), (1, 1, 17, 0), (12, 4, 16 * 5000, 5), (14, 2, 16 * 2500, 5),
    (20, 4, 16 * 2500, 5), (24, 6, 16 * 1111, 5), (6, 3, 16 * 9999, 5), (8, 4, 16 * 4096, 5),
]
for k, p, L, variant in cases:
    data = rng.integers(0, 256, (k, L), dtype=np.uint8)
    with rs.RSContext(k, p) as ctx:
        ctx.set_tuning(variant=variant)
        d_in = ctx.dev_alloc(k * L)
        d_out = ctx.dev_alloc(p * L)
        ctx.h2d(d_in, data)
        ctx.encode(d_in, L, d_out, L, L)
        ctx.sync()
        got = np.zeros((p, L), np.uint8)
        ctx.d2h(got, d_out)
        assert (got == orc.matrix_mul(orc.gen_encoding_matrix(p, k), data)).all(), (k, p, L, variant)
        # decode: erase the first min(p, k) natives
        e = min(p, k)
        order = list(range(k, k + e)) + list(range(e, k))
        try:
            ctx.set_survivors(order)
        except rs.SingularError:
            ctx.dev_free(d_in); ctx.dev_free(d_out)
            print("singular (expected for some sets)", k, p)
            continue
        surv = np.ascontiguousarray(np.concatenate([data, got], axis=0)[order])
        d_s = ctx.dev_alloc(k * L)
        d_r = ctx.dev_alloc(e * L)
        ctx.h2d(d_s, surv)
        ctx.decode_erased(d_s, L, d_r, L, L)
        ctx.sync()
        rec = np.zeros((e, L), np.uint8)
        ctx.d2h(rec, d_r)
        assert (rec == data[:e]).all(), ("decode", k, p, L)
        # host pipeline
        out2 = np.zeros((p, L), np.uint8)
        ctx.encode_host(data, L, out2, L, L, 2)
        assert (out2 == got).all()
        for ptr in (d_in, d_out, d_s, d_r):
            ctx.dev_free(ptr)
    print("ok", k, p, L, variant, flush=True)
# no-alloc entry point
with rs.RSContext(4, 2) as ctx:
    k, p, L = 10, 4, 16 * 1234 + 7
    data = rng.integers(0, 256, (k, L), dtype=np.uint8)
    d_in = ctx.dev_alloc(k * L); d_out = ctx.dev_alloc(p * L)
    ctx.h2d(d_in, data)
    rs.encode_noalloc(k, p, d_in, L, d_out, L, L)
    ctx.sync()
    got = np.zeros((p, L), np.uint8); ctx.d2h(got, d_out)
    assert (got == orc.matrix_mul(orc.gen_encoding_matrix(p, k), data)).all()
    ctx.dev_free(d_in); ctx.dev_free(d_out)
print("sanitize_smoke: all ok")
