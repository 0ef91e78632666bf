import importlib, sys, os
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
rs = importlib.import_module("gpu-rscode_b200")
from oracle_lib import Oracle
orc = Oracle()
for k, p, L in [(4, 2, 16 * 1000), (4, 2, 16 * 40000), (8, 4, 16 * 5000), (10, 4, 16 * 5000), (16, 4, 16 * 3000)]:
    rng = np.random.default_rng(1)
    data = rng.integers(0, 256, (k, L), dtype=np.uint8)
    want = orc.matrix_mul(orc.gen_encoding_matrix(p, k), data)
    with rs.RSContext(k, p) as ctx:
        ctx.set_tuning(variant=5)
        d = torch.from_numpy(data).cuda()
        out = torch.zeros((p, L), dtype=torch.uint8, device="cuda")
        ctx.encode(d, L, out, L, L)
        torch.cuda.synchronize()
        got = out.cpu().numpy()
    bad = np.argwhere(got != want)
    print(k, p, L, "mismatches", len(bad), "of", got.size)
    if len(bad):
        cols = np.unique(bad[:, 1])
        print("  rows", np.unique(bad[:, 0]), "first cols", cols[:20], "col%16", np.unique(cols % 16), "vec idx range", cols.min() // 16, cols.max() // 16)
        # which input byte values are involved? test single-fragment contributions
        for j in range(k):
            d1 = np.zeros_like(data); d1[j] = data[j]
            w1 = orc.matrix_mul(orc.gen_encoding_matrix(p, k), d1)
            with rs.RSContext(k, p) as ctx:
                ctx.set_tuning(variant=5)
                dd = torch.from_numpy(d1).cuda(); o = torch.zeros((p, L), dtype=torch.uint8, device="cuda")
                ctx.encode(dd, L, o, L, L); torch.cuda.synchronize()
                g1 = o.cpu().numpy()
            b1 = np.argwhere(g1 != w1)
            vals = np.unique(d1[j][np.unique(b1[:, 1])]) if len(b1) else []
            print("   fragment", j, "mismatches", len(b1), "bad input values (first 16)", list(vals[:16]), "n", len(vals))
