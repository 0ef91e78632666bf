"""GPU parity tests: the CUDA path, called through the C ABI (ctypes), against the CPU oracle on the
same seeded inputs and against the golden fixtures made from the reference's own cpu-rs.c.
Bit-exact everywhere: the arithmetic is integer GF(2^8)."""
import hashlib
import json
import os

import numpy as np
import pytest

from oracle_lib import Oracle
from rsdata import pattern_bytes, splitmix64_bytes

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "golden.json")))


def sha(b):
    return hashlib.sha256(bytes(b)).hexdigest()


@pytest.fixture(scope="module")
def torch():
    import torch as t
    if not t.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return t


@pytest.fixture(scope="module")
def rs(torch):
    from gpu_lib import rs as m
    m.lib()  # raises if librsgpu.so is missing: no fallback
    return m


@pytest.fixture(scope="module")
def orc():
    return Oracle()


def gf_rank(orc, M):
    """Rank over GF(2^8) by plain row reduction with the oracle's field ops (test-side check only)."""
    A = [list(map(int, r)) for r in M]
    n, m, rank = len(A), len(A[0]), 0
    for c in range(m):
        piv = next((r for r in range(rank, n) if A[r][c]), None)
        if piv is None:
            continue
        A[rank], A[piv] = A[piv], A[rank]
        inv = orc.gf_div(1, A[rank][c])
        A[rank] = [orc.gf_mul(inv, x) for x in A[rank]]
        for r in range(n):
            if r != rank and A[r][c]:
                f = A[r][c]
                A[r] = [x ^ orc.gf_mul(f, y) for x, y in zip(A[r], A[rank])]
        rank += 1
    return rank


def to_dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def gpu_encode(torch, ctx, data, pitch=None, out_pitch=None, in_off=0, out_off=0):
    """data: (k, L) uint8 host array -> (p, L) parity computed on the GPU via rsgpu_encode."""
    k, L = data.shape
    pitch = pitch or L
    out_pitch = out_pitch or L
    buf = torch.zeros(k * pitch + in_off + 64, dtype=torch.uint8, device="cuda")
    view = buf[in_off:in_off + k * pitch].view(k, pitch)
    view[:, :L] = to_dev(torch, data)
    out = torch.full((ctx.p * out_pitch + out_off + 64,), 0xEE, dtype=torch.uint8, device="cuda")
    ctx.encode(buf.data_ptr() + in_off, pitch, out.data_ptr() + out_off, out_pitch, L,
               torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    res = out[out_off:out_off + ctx.p * out_pitch].view(ctx.p, out_pitch)
    if out_pitch > L:  # bytes between rows must be untouched
        assert bool((res[:, L:] == 0xEE).all())
    return res[:, :L].cpu().numpy()


# ---- matrices -----------------------------------------------------------------------------------

@pytest.mark.parametrize("k,p", [(4, 2), (10, 4), (16, 4), (32, 8), (5, 3), (20, 12), (64, 2), (128, 3), (255, 1), (1, 1), (17, 0)])
def test_total_matrix_vs_oracle(rs, orc, k, p):
    with rs.RSContext(k, p) as ctx:
        T = ctx.total_matrix()
    assert (T == orc.gen_total_encoding_matrix(k, p)).all()
    key = f"{k},{p}"
    if key in GOLD["vandermonde"]:
        assert T[k:].tobytes().hex() == GOLD["vandermonde"][key]


def test_create_rejects_bad_arguments(rs):
    for k, p in [(0, 2), (256, 1), (4, -1), (4, 256)]:
        with pytest.raises(rs.RSGPUError) as ei:
            rs.RSContext(k, p)
        assert ei.value.code == rs.RSGPU_EINVAL
    with pytest.raises(rs.RSGPUError):
        rs.RSContext(4, 2, device=99)


# ---- encode -------------------------------------------------------------------------------------

ENC_SHAPES = [
    (4, 2, 1), (4, 2, 15), (4, 2, 16), (4, 2, 17), (10, 4, 8), (10, 4, 4096), (10, 4, 100_003),
    (16, 4, 65_536 + 48), (32, 8, 30_011), (5, 3, 4097), (6, 6, 70_001), (20, 12, 5_021), (3, 1, 777),
    (11, 4, 9_999), (13, 5, 12_345), (22, 7, 3_333), (1, 4, 1000), (64, 2, 2048), (255, 8, 512), (9, 16, 1600),
]


@pytest.mark.parametrize("k,p,L", ENC_SHAPES)
def test_encode_vs_oracle(rs, orc, torch, k, p, L):
    data = splitmix64_bytes(1000 + k * 31 + p, k * L).reshape(k, L)
    with rs.RSContext(k, p) as ctx:
        got = gpu_encode(torch, ctx, data)
    want = orc.matrix_mul(orc.gen_encoding_matrix(p, k), data)
    assert (got == want).all()


@pytest.mark.parametrize("variant", [0, 1, 5, 6, 7, 9])  # auto, nibble+PRMT, byte tables, nibble+DP4A, nibble fed by TMA, multi-pass
@pytest.mark.parametrize("k,p,L", [(10, 4, 200_000), (32, 8, 50_000), (11, 3, 40_001), (13, 5, 160_000), (19, 2, 64_000), (23, 8, 32_000),
                                     (16, 4, 70_000 * 16), (5, 2, 4096 * 16 + 48), (4, 8, 100_000)])
def test_encode_variants_vs_oracle(rs, orc, torch, variant, k, p, L):
    data = splitmix64_bytes(77 + variant, k * L).reshape(k, L)
    with rs.RSContext(k, p) as ctx:
        ctx.set_tuning(variant=variant)
        got = gpu_encode(torch, ctx, data)
    assert (got == orc.matrix_mul(orc.gen_encoding_matrix(p, k), data)).all()


@pytest.mark.parametrize("k", list(range(2, 33)))
@pytest.mark.parametrize("p", [1, 4, 5, 8])
def test_encode_bank_replicated_kernel_vs_oracle(rs, orc, torch, k, p):
    """variant 5 forces the byte-table kernel (fat8 / fat7 / nibble mix depends on k) at a small size."""
    L = 150_000 + 16 * k + 3
    data = splitmix64_bytes(500 + k * 7 + p, k * L).reshape(k, L)
    data[:, :512] = np.arange(512, dtype=np.uint8)[None, :]          # every byte value, bit 7 included
    with rs.RSContext(k, p) as ctx:
        ctx.set_tuning(variant=5)
        pitch = (L + 255) // 256 * 256                                # 16-byte aligned rows: the vector kernel runs
        got = gpu_encode(torch, ctx, data, pitch=pitch, out_pitch=pitch)
        n_launch = ctx.launch_count()
    assert (got == orc.matrix_mul(orc.gen_encoding_matrix(p, k), data)).all()
    assert n_launch > 0


@pytest.mark.parametrize("k,p", [(15, 4), (25, 4), (28, 4), (29, 3), (32, 8), (33, 1), (40, 5), (64, 4), (100, 6), (128, 2), (255, 4), (42, 13)])
def test_encode_multipass_kernel_vs_oracle(rs, orc, torch, k, p):
    """variant 9 forces the multi-pass byte-table kernel (fragments taken 14 at a time, partial parities in registers
    across the table rebuilds): one block step exactly, several steps plus a remainder that goes down the ordinary
    path, a ragged tail; last pass short (k % 14 != 0), one full pass plus one fragment, k = 255."""
    V = orc.gen_encoding_matrix(p, k)
    with rs.RSContext(k, p) as ctx:
        ctx.set_tuning(variant=9)
        for L in (24_576, 3 * 24_576 + 16 * 37 + 5, 200_000):
            data = splitmix64_bytes(7000 + k * 3 + p + L % 13, k * L).reshape(k, L)
            data[:, :256] = np.arange(256, dtype=np.uint8)[None, :]
            pitch = (L + 255) // 256 * 256
            n0 = ctx.launch_count()
            got = gpu_encode(torch, ctx, data, pitch=pitch, out_pitch=pitch)
            assert ctx.launch_count() > n0
            bad = np.argwhere(got != orc.matrix_mul(V, data))
            assert bad.size == 0, f"L={L}: first mismatch at {bad[0]} of {len(bad)}"


def test_encode_multipass_kernel_long_stripe(rs, orc, torch):
    """The multi-pass kernel (variant 9) on a long stripe, many block steps per SM: oracle on sampled windows, and the
    same bytes as the nibble kernel (variant 6)."""
    k, p, L = 40, 4, (8 << 20) + 16 * 100 + 7
    pitch = (L + 255) // 256 * 256
    st = torch.cuda.current_stream().cuda_stream
    data = torch.randint(0, 256, (k, pitch), dtype=torch.uint8, device="cuda")
    out = torch.full((p, pitch), 0xEE, dtype=torch.uint8, device="cuda")
    out6 = torch.full((p, pitch), 0xEE, dtype=torch.uint8, device="cuda")
    V = orc.gen_encoding_matrix(p, k)
    with rs.RSContext(k, p) as ctx:
        ctx.set_tuning(variant=9)
        ctx.encode(data, pitch, out, pitch, L, st)
        ctx.set_tuning(variant=6)
        ctx.encode(data, pitch, out6, pitch, L, st)
        torch.cuda.synchronize()
    assert torch.equal(out, out6)
    assert bool((out[:, L:] == 0xEE).all())
    rng = np.random.default_rng(5)
    for c0 in [0, L - 50_000] + [int(x) for x in rng.integers(0, L - 50_000, 6)]:
        want = orc.matrix_mul(V, data[:, c0:c0 + 50_000].cpu().numpy())
        assert (out[:, c0:c0 + 50_000].cpu().numpy() == want).all(), f"window at {c0}"


@pytest.mark.parametrize("k,p", [(10, 5), (15, 5), (4, 5), (7, 13), (25, 5), (40, 5), (16, 21)])
def test_encode_with_peeled_all_ones_row_vs_oracle(rs, orc, torch, k, p):
    """Five parity rows (p % 8 == 5): row 0 of the generator is all ones ((j+1)^0, src/matrix.cu:758) and is
    computed as the XOR of the fragments beside four looked-up rows on 32-bit entries.  Every dispatch path must give
    the oracle's bytes: byte kernel (short / unaligned short), nibble kernel, byte tables (long and forced), the
    realigning kernel, the batch kernel, per-row destinations, the host pipeline; a caller-supplied matrix with the
    same rows peels too, a Cauchy generator must not."""
    st = torch.cuda.current_stream().cuda_stream
    V = orc.gen_encoding_matrix(p, k)
    assert (V[0] == 1).all()

    def check(ctx, L, pitch=None, in_off=0, out_off=0, variant=0, Vm=V):
        data = splitmix64_bytes(3000 + k * 13 + L % 1009, k * L).reshape(k, L)
        data[:, :min(L, 256)] = np.arange(256, dtype=np.uint8)[None, :min(L, 256)]
        ctx.set_tuning(variant=variant)
        got = gpu_encode(torch, ctx, data, pitch=pitch, out_pitch=pitch, in_off=in_off, out_off=out_off)
        bad = np.argwhere(got != orc.matrix_mul(Vm, data))
        assert bad.size == 0, f"L={L} pitch={pitch} variant={variant}: first mismatch at {bad[0]}"

    long_ok = k <= 32
    with rs.RSContext(k, p) as ctx:
        check(ctx, 1000)                                        # one launch of the byte kernel
        check(ctx, 70_001, pitch=70_016)                        # nibble kernel + byte tail
        check(ctx, 40_000, in_off=3, out_off=9)                 # unaligned short: byte kernel
        check(ctx, 300_000, pitch=300_032, variant=5)           # byte tables forced (k <= 32) or the nibble kernel
        check(ctx, 300_000, pitch=300_032, variant=1)           # PRMT variant has no peeled instance: DP4A one runs
        check(ctx, (3 << 20) + 48, pitch=(3 << 20) + 256)       # 2-8 MiB: the TMA kernel steps aside
        if long_ok and k <= 16:
            check(ctx, (8 << 20) + 16)                          # long stripe: byte tables by default
        check(ctx, (1 << 20) + 5, in_off=1, out_off=2)          # realigning kernel (k <= 32), else byte kernel
        # batch: 6 stripes, stripe-major, ragged length
        Ls, n = 70_000 + 5, 6
        sp_in, sp_out = k * 70_016, p * 70_016
        data = splitmix64_bytes(4000 + k, n * k * Ls).reshape(n, k, Ls)
        buf = torch.zeros(n * sp_in, dtype=torch.uint8, device="cuda")
        buf.view(n, k, 70_016)[:, :, :Ls] = to_dev(torch, data)
        out = torch.full((n * sp_out,), 0xEE, dtype=torch.uint8, device="cuda")
        ctx.set_tuning(variant=5)
        ctx.encode_batch(buf, 70_016, sp_in, out, 70_016, sp_out, Ls, n, st)
        torch.cuda.synchronize()
        got = out.view(n, p, 70_016)[:, :, :Ls].cpu().numpy()
        for s_i in range(n):
            assert (got[s_i] == orc.matrix_mul(V, data[s_i])).all(), f"batch stripe {s_i}"
        assert bool((out.view(n, p, 70_016)[:, :, Ls:] == 0xEE).all())
        ctx.set_tuning(variant=0)
        # one destination pointer per row (same device)
        L = 100_000 + 16 * 3 + 5
        data = splitmix64_bytes(5000 + k, k * L).reshape(k, L)
        rows = [torch.full((L + 64,), 0xEE, dtype=torch.uint8, device="cuda") for _ in range(p)]
        ctx.encode_rows(to_dev(torch, data), L, rows, L, st)
        torch.cuda.synchronize()
        want = orc.matrix_mul(V, data)
        for i, r in enumerate(rows):
            assert (r[:L].cpu().numpy() == want[i]).all() and bool((r[L:] == 0xEE).all()), f"row {i}"
        # host pipeline
        L = 600_003
        hd = splitmix64_bytes(6000 + k, k * L).reshape(k, L)
        hp = np.zeros((p, L), np.uint8)
        ctx.encode_host(hd, L, hp, L, L, 2)
        assert (hp == orc.matrix_mul(V, hd)).all()
    # the same matrix handed in by the caller peels as well (row k of the total matrix is all ones) ...
    T = orc.gen_total_encoding_matrix(k, p)
    with rs.RSContext(k, p, total_matrix=T) as ctx:
        check(ctx, 300_000, pitch=300_032, variant=5)
    # ... a total matrix whose first parity row is NOT all ones must not (rows swapped: the all-ones row comes second)
    T2 = T.copy()
    T2[[k, k + 1]] = T2[[k + 1, k]]
    with rs.RSContext(k, p, total_matrix=T2) as ctx:
        check(ctx, 300_000, pitch=300_032, variant=5, Vm=T2[k:])
    if k + p <= 256:
        with rs.RSContext(k, p, generator="cauchy") as ctx:
            C_ = ctx.generator()
            check(ctx, 300_000, pitch=300_032, variant=5, Vm=C_)


@pytest.mark.parametrize("k", [17, 18, 20, 21, 23, 24, 25, 27, 28, 29, 31, 32])
@pytest.mark.parametrize("p,L", [(4, 150_003), (3, 16 * 40), (1, 16 * 100_000), (2, 16 * 7 + 5), (4, 16 * 777_777)])
def test_encode_cluster_kernel_vs_oracle(rs, orc, torch, k, p, L):
    """variant 8 forces the 2-CTA cluster kernel (fragments split over the pair, partial parities exchanged
    through distributed shared memory): ragged tails, stripes shorter than one warp, long stripes."""
    if L > 16 * 200_000 and k not in (17, 24, 32):
        pytest.skip("long stripe only for three k")
    data = splitmix64_bytes(800 + k * 5 + p, k * L).reshape(k, L)
    data[:, :min(L, 512)] = np.arange(512, dtype=np.uint8)[None, :min(L, 512)]
    with rs.RSContext(k, p) as ctx:
        ctx.set_tuning(variant=8)
        pitch = (L + 255) // 256 * 256
        got = gpu_encode(torch, ctx, data, pitch=pitch, out_pitch=pitch)
    assert (got == orc.matrix_mul(orc.gen_encoding_matrix(p, k), data)).all()


def test_encode_pitch_and_misalignment(rs, orc, torch):
    k, p, L = 10, 4, 10_007
    data = splitmix64_bytes(5, k * L).reshape(k, L)
    want = orc.matrix_mul(orc.gen_encoding_matrix(p, k), data)
    with rs.RSContext(k, p) as ctx:
        assert (gpu_encode(torch, ctx, data, pitch=10_240, out_pitch=12_288) == want).all()   # aligned pitches
        assert (gpu_encode(torch, ctx, data, pitch=10_009, out_pitch=10_011) == want).all()   # odd pitches -> byte path
        assert (gpu_encode(torch, ctx, data, pitch=10_240, out_pitch=10_240, in_off=3) == want).all()
        assert (gpu_encode(torch, ctx, data, pitch=10_240, out_pitch=10_240, out_off=5) == want).all()


UNALIGNED_CASES = [
    # k, p, L, in_pitch - L, out_pitch - L, in_off, out_off
    (10, 4, (1 << 20) + 3, 0, 0, 0, 0),            # pitch = L odd: the layout of a file's fragments held back to back
    (10, 4, (3 << 20) + 7, 0, 0, 5, 11),           # bases unaligned too
    (10, 4, (1 << 20), 16, 16, 8, 4),              # 16-byte pitches, only the bases are off (every row the same shift)
    (10, 4, (2 << 20) + 1, 1, 2, 15, 1),
    (4, 2, (5 << 20) + 9, 0, 0, 1, 2),             # all byte tables (fat8)
    (16, 4, (1 << 20) + 5, 0, 0, 3, 9),            # transposed bit-7 accumulators
    (15, 5, (1 << 20) + 13, 0, 0, 7, 7),           # padded shape, five parity rows
    (32, 8, (1 << 20) + 11, 0, 0, 2, 6),           # 64-bit entries, two row halves
    (24, 3, (1 << 20) + 6, 0, 0, 0, 12),
    (7, 3, (1 << 20) + 2, 4, 8, 4, 8),             # word-aligned shifts only
    (12, 9, (1 << 20) + 3, 0, 0, 9, 3),            # two row groups (8 + 1)
    (9, 4, (1 << 20) + 3, 0, 0, 6, 10),            # odd batch count (3 x 3)
    (5, 3, 1_048_576 + 480 * 7 + 1, 0, 0, 13, 14),
    # k without a byte-table shape: the same realignment on split-nibble tables
    (25, 4, (1 << 20) + 3, 0, 0, 1, 2),
    (28, 3, (1 << 20) + 9, 0, 0, 0, 0),
    (40, 5, (1 << 20) + 1, 0, 0, 5, 5),            # peeled XOR row
    (64, 8, (1 << 20) + 5, 0, 0, 3, 7),            # 64-bit entries
    (100, 2, (1 << 20) + 7, 2, 6, 9, 9),
    (255, 1, (1 << 20) + 11, 0, 0, 15, 0),
]


@pytest.mark.parametrize("k,p,L,dpi,dpo,in_off,out_off", UNALIGNED_CASES)
def test_encode_unaligned_rows_run_the_vector_kernel(rs, orc, torch, k, p, L, dpi, dpo, in_off, out_off):
    """Rows that are not 16-byte aligned (odd pitch and/or odd base) must stay on the byte-table kernel --
    aligned 128-bit accesses realigned in registers -- with the byte kernel only for the head and the tail; every
    byte equals the oracle's and nothing outside the output rows is touched."""
    data = splitmix64_bytes(900 + k + L % 97, k * L).reshape(k, L)
    want = orc.matrix_mul(orc.gen_encoding_matrix(p, k), data)
    pi, po = L + dpi, L + dpo
    buf = torch.zeros(k * pi + in_off + 64, dtype=torch.uint8, device="cuda")
    buf[in_off:in_off + k * pi].view(k, pi)[:, :L] = to_dev(torch, data)
    out = torch.full((p * po + out_off + 64,), 0xEE, dtype=torch.uint8, device="cuda")
    with rs.RSContext(k, p) as ctx:
        n0 = ctx.launch_count()
        ctx.encode(buf.data_ptr() + in_off, pi, out.data_ptr() + out_off, po, L, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        groups = (p + 7) // 8
        assert ctx.launch_count() - n0 == 3 * groups, "expected vector kernel + head + tail per row group"
        # the byte kernel alone (variant 6) gives the same bytes
        out6 = torch.full_like(out, 0xEE)
        ctx.set_tuning(variant=6)
        n0 = ctx.launch_count()
        ctx.encode(buf.data_ptr() + in_off, pi, out6.data_ptr() + out_off, po, L, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert ctx.launch_count() - n0 == groups
    res = out[out_off:out_off + p * po].view(p, po)
    got = res[:, :L].cpu().numpy()
    bad = np.argwhere(got != want)
    assert bad.size == 0, f"first mismatch at row/byte {bad[0]} of {len(bad)}"
    assert bool((out[:out_off] == 0xEE).all()) and bool((out[out_off + p * po:] == 0xEE).all())
    if po > L:
        assert bool((res[:-1, L:] == 0xEE).all())
    assert torch.equal(out, out6)


def test_encode_unaligned_large_stripe_and_decode(rs, orc, torch):
    """A long stripe with pitch = L odd (RS(10,14), 64 MiB + 3 per fragment): encode and 4-erasure decode through the
    realigning kernel; oracle on sampled windows of every row, full round trip on the device."""
    k, p, L = 10, 4, (64 << 20) + 3
    st = torch.cuda.current_stream().cuda_stream
    data = torch.randint(0, 256, (k * L + 1,), dtype=torch.uint8, device="cuda")[1:]      # base 1 byte off
    par = torch.empty(p * L + 7, dtype=torch.uint8, device="cuda")[7:]
    V = orc.gen_encoding_matrix(p, k)
    with rs.RSContext(k, p) as ctx:
        ctx.encode(data.data_ptr(), L, par.data_ptr(), L, L, st)
        torch.cuda.synchronize()
        d2, p2 = data.view(k, L), par.view(p, L)
        rng = np.random.default_rng(3)
        for c0 in [0, L - 70_000] + [int(x) for x in rng.integers(0, L - 70_000, 12)]:
            want = orc.matrix_mul(V, d2[:, c0:c0 + 70_000].cpu().numpy())
            assert (p2[:, c0:c0 + 70_000].cpu().numpy() == want).all(), f"window at {c0}"
        order = [10, 11, 12, 13, 4, 5, 6, 7, 8, 9]
        ctx.set_survivors(order)
        surv = torch.cat([p2, d2[4:]], dim=0).contiguous().view(-1)                      # k x L, pitch L (odd)
        out = torch.empty(4 * L + 5, dtype=torch.uint8, device="cuda")[5:]
        ctx.decode_erased(surv.data_ptr(), L, out.data_ptr(), L, L, st)
        torch.cuda.synchronize()
        assert torch.equal(out.view(4, L), d2[:4])


def test_encode_overwrites_and_zero_len(rs, torch):
    with rs.RSContext(4, 2) as ctx:
        d = torch.zeros(4 * 64, dtype=torch.uint8, device="cuda")
        out = torch.full((2 * 64,), 0x55, dtype=torch.uint8, device="cuda")
        ctx.encode(d, 64, out, 64, 0)          # empty input: nothing written, no error
        torch.cuda.synchronize()
        assert bool((out == 0x55).all())
        ctx.encode(d, 64, out, 64, 64)         # zero data -> zero parity even though out was dirty
        torch.cuda.synchronize()
        assert bool((out == 0).all())
        with pytest.raises(rs.RSGPUError):
            ctx.encode(d, 32, out, 64, 64)     # pitch < len


@pytest.mark.parametrize("ci", range(len(GOLD["encode"])))
def test_encode_golden_fragments(rs, torch, ci):
    """Same file split as cpu-rs.c encode_file (zero-padded last native), parity fragments hashed."""
    spec = GOLD["encode"][ci]
    k, p, L = spec["k"], spec["p"], spec["chunk"]
    raw = pattern_bytes(spec["mul"], spec["add"], spec["nbytes"]) if spec["gen"] == "pattern" \
        else splitmix64_bytes(spec["seed"], spec["nbytes"])
    data = np.zeros(k * L, np.uint8)
    data[:spec["nbytes"]] = raw
    data = data.reshape(k, L)
    with rs.RSContext(k, p) as ctx:
        got = gpu_encode(torch, ctx, data)
    for i in range(k):
        assert sha(data[i].tobytes()) == spec["fragment_sha256"][i]
    for i in range(p):
        assert sha(got[i].tobytes()) == spec["fragment_sha256"][k + i], f"parity {i}"
        if "fragment_hex" in spec:
            assert got[i].tobytes().hex() == spec["fragment_hex"][k + i]


# ---- inverse ------------------------------------------------------------------------------------

@pytest.mark.parametrize("case", GOLD["inverse"] + GOLD["inverse_natural_rs46"],
                         ids=lambda c: f"k{c['k']}-{'_'.join(map(str, c['order'][:5]))}")
def test_inverse_vs_golden(rs, orc, case):
    k, p = case["k"], case["p"]
    with rs.RSContext(k, p) as ctx:
        ctx.set_survivors(case["order"])
        inv = ctx.decode_matrix()
    if case.get("is_true_inverse", True):
        assert inv.tobytes().hex() == case["inverse"]      # unique inverse: same bytes as the reference
    sub = orc.gen_total_encoding_matrix(k, p)[case["order"]]
    assert (orc.matrix_mul(sub, inv) == np.eye(k, dtype=np.uint8)).all()


def test_inverse_kats_from_reference_docs(rs):
    with rs.RSContext(4, 2) as ctx:
        M = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [1, 2, 3, 4], [1, 1, 1, 1]], np.uint8)   # doc/slide.tex:501-557
        assert ctx.invert_matrix(M).tolist() == [[1, 0, 0, 0], [0, 1, 0, 0], [104, 187, 186, 210], [105, 186, 186, 211]]
        M = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [1, 1, 1, 1], [1, 2, 3, 4]], np.uint8)   # src/decode-gj.cu:1282,1453
        assert ctx.invert_matrix(M).tolist() == [[1, 0, 0, 0], [0, 1, 0, 0], [104, 187, 210, 186], [105, 186, 211, 186]]


def test_inverse_random_orders_true_inverse(rs, orc):
    """Any survivor ORDER, also those where the reference's column swap goes wrong: A * inv == I; and
    bit-identical to the oracle wherever the oracle reports the reference-correct domain."""
    rng = np.random.default_rng(3)
    for k, p in [(4, 2), (10, 4), (16, 4), (32, 8)]:
        T = orc.gen_total_encoding_matrix(k, p)
        with rs.RSContext(k, p) as ctx:
            n_ok = n_same = 0
            for _ in range(60):
                order = rng.permutation(k + p)[:k].tolist()
                try:
                    ctx.set_survivors(order)
                except rs.SingularError:
                    assert gf_rank(orc, T[order]) < k      # independently confirmed singular
                    continue
                inv = ctx.decode_matrix()
                assert (orc.matrix_mul(T[order], inv) == np.eye(k, dtype=np.uint8)).all()
                n_ok += 1
                ref_inv, st = orc.invert_matrix(T[order])
                if st == 0:
                    assert (inv == ref_inv).all()
                    n_same += 1
            assert n_ok >= 40


def test_inverse_large_random_matrices(rs, orc):
    rng = np.random.default_rng(11)
    with rs.RSContext(4, 2) as ctx:
        for n in (1, 2, 7, 33, 100, 255):
            M = rng.integers(0, 256, (n, n), dtype=np.uint8)
            try:
                inv = ctx.invert_matrix(M)
            except rs.SingularError:
                continue
            assert (orc.matrix_mul(M, inv) == np.eye(n, dtype=np.uint8)).all()


def test_singular_survivor_sets_are_reported(rs):
    # [I;V] is not MDS; RS(10,14) erasing natives {0,1,2} and parity 12 is singular
    k, p = 10, 4
    with rs.RSContext(k, p) as ctx:
        for erased in [(0, 1, 2, 12), (0, 3, 4, 12), (4, 5, 9, 11)]:
            surv = [i for i in range(k + p) if i not in erased]
            with pytest.raises(rs.SingularError):
                ctx.set_survivors(surv)
        ctx.set_survivors([i for i in range(14) if i not in (0, 1, 2, 3)])   # fine again afterwards
    with rs.RSContext(4, 2) as ctx:
        with pytest.raises(rs.SingularError):
            ctx.invert_matrix(np.array([[1, 2, 0, 0], [2, 4, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1]], np.uint8))
        with pytest.raises(rs.RSGPUError):
            ctx.set_survivors([0, 1, 2, 9])      # index out of range
        d = ctx.dev_alloc(64)
        with pytest.raises(rs.RSGPUError) as ei:  # decode before a successful set_survivors
            rs.RSContext(4, 2).decode(d, 16, d, 16, 16)
        assert ei.value.code == rs.RSGPU_ESTATE
        ctx.dev_free(d)


# ---- decode -------------------------------------------------------------------------------------

DEC_CASES = [
    (4, 2, [0, 1, 4, 5], 16_777),
    (4, 2, [2, 3, 4, 5], 4_099),            # the unit-test.sh order: reference decodes this WRONG, we must not
    (4, 2, [0, 1, 3, 5], 1_000),            # ditto
    (10, 4, [10, 11, 12, 13, 4, 5, 6, 7, 8, 9], 100_003),
    (10, 4, [0, 1, 2, 10, 4, 5, 11, 7, 8, 13], 33_333),
    (16, 4, [16, 17, 18, 19] + list(range(4, 16)), 20_480),
    (32, 8, list(range(32, 40)) + list(range(8, 32)), 9_001),
    (5, 3, [7, 6, 5, 4, 3], 2_222),
    (10, 4, list(range(10)), 5_000),        # nothing erased
    (200, 20, [200, 201, 202] + list(range(3, 200)), 4_096),    # large k: 1024-thread set-up kernel, 25 row groups
    (255, 1, [255] + list(range(1, 255)), 2_048),
]


@pytest.mark.parametrize("k,p,order,L", DEC_CASES, ids=lambda v: str(v)[:24])
def test_decode_roundtrip_and_oracle(rs, orc, torch, k, p, order, L):
    data = splitmix64_bytes(4242 + k, k * L).reshape(k, L)
    V = orc.gen_encoding_matrix(p, k)
    frags = np.concatenate([data, orc.matrix_mul(V, data)], axis=0)       # oracle-encoded stripe
    surv = np.ascontiguousarray(frags[order])
    with rs.RSContext(k, p) as ctx:
        ctx.set_survivors(order)
        d_surv = to_dev(torch, surv)
        out = torch.empty((k, L), dtype=torch.uint8, device="cuda")
        ctx.decode(d_surv, L, out, L, L, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert (out.cpu().numpy() == data).all()
        # reference-semantic result == oracle's matrix_mul with the oracle's inverse where that is right
        ref_inv, st = orc.invert_matrix(orc.gen_total_encoding_matrix(k, p)[order])
        if st == 0:
            assert (orc.matrix_mul(ref_inv, surv) == out.cpu().numpy()).all()
        erased = ctx.erased()
        assert erased == sorted(set(range(k)) - set(i for i in order if i < k))
        if erased:
            out_e = torch.empty((len(erased), L), dtype=torch.uint8, device="cuda")
            ctx.decode_erased(d_surv, L, out_e, L, L, torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
            assert (out_e.cpu().numpy() == data[erased]).all()


def test_set_survivors_while_previous_decode_is_in_flight(rs, torch):
    """rsgpu_set_survivors for the NEXT survivor set may be called while decodes of the previous set are
    still queued: each decode must finish with the tables it was launched with (two plan sets, event
    ordered).  Long stripes keep the kernels busy while the host races ahead."""
    k, p, L = 10, 4, 48 << 20
    stream = torch.cuda.current_stream().cuda_stream
    data = torch.randint(0, 256, (k, L), dtype=torch.uint8, device="cuda")
    orders = [[10, 11, 12, 13, 4, 5, 6, 7, 8, 9], [0, 1, 2, 3, 4, 5, 10, 11, 12, 13], [13, 1, 12, 3, 11, 5, 10, 7, 8, 9],
              list(range(10)), [0, 11, 2, 13, 4, 5, 6, 7, 12, 9]]
    with rs.RSContext(k, p) as ctx:
        parity = torch.empty((p, L), dtype=torch.uint8, device="cuda")
        ctx.encode(data.data_ptr(), L, parity.data_ptr(), L, L, stream)
        frags = torch.cat([data, parity])
        survs = [frags[torch.tensor(o, device="cuda")].contiguous() for o in orders]
        outs = [torch.empty((k, L), dtype=torch.uint8, device="cuda") for _ in range(3 * len(orders))]
        torch.cuda.synchronize()
        for i, out in enumerate(outs):            # no synchronisation between the steps
            o = i % len(orders)
            ctx.set_survivors(orders[o])
            if i % 2:
                ctx.decode(survs[o].data_ptr(), L, out.data_ptr(), L, L, stream)
            else:
                out.copy_(data)                   # decode_erased writes only the lost natives
                er = ctx.erased()
                if er:
                    tmp = torch.empty((len(er), L), dtype=torch.uint8, device="cuda")
                    ctx.decode_erased(survs[o].data_ptr(), L, tmp.data_ptr(), L, L, stream)
                    out[torch.tensor(er, device="cuda")] = tmp
        torch.cuda.synchronize()
        for i, out in enumerate(outs):
            assert torch.equal(out, data), f"step {i}"


def test_decode_on_several_streams_right_after_set_survivors(rs, torch):
    """One context, several streams: the tables of the full k x k plan are built on first use, so a decode on a
    SECOND stream must wait for that build (queued by the first), and rsgpu_set_survivors must wait for the
    decodes of every stream before it rewrites a table set -- not only for the stream used last.  Six streams
    (more than the tracker has slots) decode concurrently, survivor sets change without host synchronisation."""
    k, p, L = 10, 4, 24 << 20
    s0 = torch.cuda.current_stream().cuda_stream
    streams = [torch.cuda.Stream() for _ in range(6)]
    data = torch.randint(0, 256, (k, L), dtype=torch.uint8, device="cuda")
    orders = [[10, 11, 12, 13, 4, 5, 6, 7, 8, 9], [0, 1, 2, 3, 4, 5, 10, 11, 12, 13], [13, 1, 12, 3, 11, 5, 10, 7, 8, 9]]
    with rs.RSContext(k, p) as ctx:
        parity = torch.empty((p, L), dtype=torch.uint8, device="cuda")
        ctx.encode(data.data_ptr(), L, parity.data_ptr(), L, L, s0)
        frags = torch.cat([data, parity])
        survs = [frags[torch.tensor(o, device="cuda")].contiguous() for o in orders]
        outs = [[torch.zeros((k, L), dtype=torch.uint8, device="cuda") for _ in streams] for _ in range(4)]
        torch.cuda.synchronize()
        for rnd, row in enumerate(outs):
            o = rnd % len(orders)
            ctx.set_survivors(orders[o])
            for st, out in zip(streams, row):      # first call builds the tables; the others must wait for them
                ctx.decode(survs[o].data_ptr(), L, out.data_ptr(), L, L, st.cuda_stream)
        torch.cuda.synchronize()
        for rnd, row in enumerate(outs):
            for i, out in enumerate(row):
                assert torch.equal(out, data), f"round {rnd}, stream {i}"


@pytest.mark.parametrize("k", [1, 2, 7, 10, 16, 33, 200])
def test_single_all_ones_row_is_a_plain_xor(rs, orc, torch, k):
    """One output row whose coefficients are all 1 runs the XOR kernel (no tables): the parity of RS(k, k+1), and the one
    lost native of a stripe that still has parity row 0.  The same stripe with parity row 1 instead takes the table
    kernels; all of it against the oracle, aligned and unaligned, with a ragged tail."""
    st = torch.cuda.current_stream().cuda_stream
    for L, pitch in ((100_000 + 5, 100_016), (70_001, 70_001), (5, 16)):
        data = splitmix64_bytes(9000 + k + L, k * L).reshape(k, L)
        with rs.RSContext(k, 1) as ctx:
            got = gpu_encode(torch, ctx, data, pitch=pitch, out_pitch=pitch)
        assert (got == orc.matrix_mul(orc.gen_encoding_matrix(1, k), data)).all()
    if k < 2 or k > 100:
        return
    p, L = 3, 300_000 + 16 * 3 + 7
    pitch = (L + 255) // 256 * 256
    data = splitmix64_bytes(9100 + k, k * L).reshape(k, L)
    V = orc.gen_encoding_matrix(p, k)
    frags = np.concatenate([data, orc.matrix_mul(V, data)], axis=0)
    lost = k // 2
    for par in (k, k + 1):                      # parity row 0 (all ones -> XOR), parity row 1 (tables)
        order = list(range(k))
        order[lost] = par
        surv = torch.zeros((k, pitch), dtype=torch.uint8, device="cuda")
        surv[:, :L] = to_dev(torch, frags[order])
        out = torch.full((1, pitch), 0xEE, dtype=torch.uint8, device="cuda")
        with rs.RSContext(k, p) as ctx:
            ctx.set_survivors(order)
            assert ctx.erased() == [lost]
            ctx.decode_erased(surv, pitch, out, pitch, L, st)
            torch.cuda.synchronize()
        assert (out[0, :L].cpu().numpy() == data[lost]).all(), f"parity {par}"
        assert bool((out[0, L:] == 0xEE).all())


def test_gpu_encode_then_gpu_decode_all_double_erasures_rs46(rs, torch):
    k, p, L = 4, 2, 3_001
    data = splitmix64_bytes(9, k * L).reshape(k, L)
    with rs.RSContext(k, p) as ctx:
        parity = gpu_encode(torch, ctx, data)
        frags = np.concatenate([data, parity], axis=0)
        import itertools
        for lost in itertools.combinations(range(k + p), 2):
            order = [i for i in range(k + p) if i not in lost][:k]
            ctx.set_survivors(order)
            d_surv = to_dev(torch, frags[order])
            out = torch.empty((k, L), dtype=torch.uint8, device="cuda")
            ctx.decode(d_surv, L, out, L, L)
            torch.cuda.synchronize()
            assert (out.cpu().numpy() == data).all(), lost


# ---- generic matmul, many rows (several row groups) -------------------------------------------

@pytest.mark.parametrize("rows,k,L", [(1, 10, 1000), (4, 10, 1000), (5, 10, 1000), (8, 7, 4000), (9, 16, 2049),
                                        (20, 12, 3000), (40, 32, 1111), (255, 3, 300)])
def test_generic_matmul_vs_oracle(rs, orc, torch, rows, k, L):
    rng = np.random.default_rng(rows * 100 + k)
    M = rng.integers(0, 256, (rows, k), dtype=np.uint8)
    B = rng.integers(0, 256, (k, L), dtype=np.uint8)
    with rs.RSContext(k, 1) as ctx:
        d_in = to_dev(torch, B)
        out = torch.empty((rows, L), dtype=torch.uint8, device="cuda")
        ctx.matmul(M, d_in, L, out, L, L)
        assert (out.cpu().numpy() == orc.matrix_mul(M, B)).all()


# ---- host-memory pipeline ------------------------------------------------------------------------

@pytest.mark.parametrize("k,p,L,slots", [(4, 2, 5_000_001, 3), (10, 4, 3_000_000, 4), (32, 8, 700_003, 2), (10, 4, 100, 1)])
def test_host_pipeline_vs_oracle(rs, orc, torch, k, p, L, slots):
    data = splitmix64_bytes(31 + k, k * L).reshape(k, L)
    want = orc.matrix_mul(orc.gen_encoding_matrix(p, k), data)
    pin_in = torch.from_numpy(data).pin_memory()
    pin_out = torch.empty((p, L), dtype=torch.uint8).pin_memory()
    with rs.RSContext(k, p) as ctx:
        ctx.encode_host(pin_in.data_ptr(), L, pin_out.data_ptr(), L, L, slots)
        assert (pin_out.numpy() == want).all()
        # pageable buffers work too (slower)
        out2 = np.zeros((p, L), np.uint8)
        ctx.encode_host(data, L, out2, L, L, slots)
        assert (out2 == want).all()
        # decode through the pipeline
        frags = np.concatenate([data, want], axis=0)
        order = list(range(k, k + p)) + list(range(p, k)) if p <= k else list(range(k))
        ctx.set_survivors(order)
        surv = np.ascontiguousarray(frags[order])
        back = np.zeros((k, L), np.uint8)
        ctx.decode_host(surv, L, back, L, L, slots)
        assert (back == data).all()


# ---- full-size, size-independent properties (BASELINE configs are too big for the oracle) ----------

def test_large_roundtrip_and_linearity_rs10_14(rs, torch):
    k, p = 10, 4
    L = 64 << 20                                # 64 MiB per fragment: 640 MiB in, 256 MiB parity
    g = torch.Generator(device="cuda").manual_seed(1)
    a = torch.randint(0, 256, (k, L), dtype=torch.uint8, device="cuda", generator=g)
    b = torch.randint(0, 256, (k, L), dtype=torch.uint8, device="cuda", generator=g)
    st = torch.cuda.current_stream().cuda_stream
    with rs.RSContext(k, p) as ctx:
        pa = torch.empty((p, L), dtype=torch.uint8, device="cuda")
        pb = torch.empty_like(pa)
        pab = torch.empty_like(pa)
        ctx.encode(a, L, pa, L, L, st)
        ctx.encode(b, L, pb, L, L, st)
        ctx.encode(a ^ b, L, pab, L, L, st)
        torch.cuda.synchronize()
        assert bool(((pa ^ pb) == pab).all())                       # linearity over GF(2)
        x = a[0].clone()
        for j in range(1, k):
            x ^= a[j]
        assert bool((pa[0] == x).all())                             # row 0 of V is all ones: plain XOR
        # encode -> erase 4 natives -> decode
        order = [10, 11, 12, 13, 4, 5, 6, 7, 8, 9]
        ctx.set_survivors(order)
        surv = torch.cat([pa, a[4:]], dim=0).contiguous()
        out = torch.empty((4, L), dtype=torch.uint8, device="cuda")
        ctx.decode_erased(surv, L, out, L, L, st)
        torch.cuda.synchronize()
        assert bool((out == a[:4]).all())


def test_sampled_windows_vs_oracle_large(rs, orc, torch):
    """Columns are independent: check head, tail and random 4 KiB windows of a large stripe with the oracle."""
    k, p = 16, 4
    L = (32 << 20) + 16 * 7 + 5                 # not a multiple of 16: exercises the tail kernel too
    g = torch.Generator(device="cuda").manual_seed(2)
    a = torch.randint(0, 256, (k, L), dtype=torch.uint8, device="cuda", generator=g)
    with rs.RSContext(k, p) as ctx:
        out = torch.empty((p, L), dtype=torch.uint8, device="cuda")
        ctx.encode(a, L, out, L, L, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
    V = orc.gen_encoding_matrix(p, k)
    rng = np.random.default_rng(4)
    starts = [0, L - 4096] + [int(s) for s in rng.integers(0, L - 4096, 30)]
    for s in starts:
        win = a[:, s:s + 4096].cpu().numpy()
        assert (out[:, s:s + 4096].cpu().numpy() == orc.matrix_mul(V, win)).all(), s


# ---- repair of arbitrary fragments, Cauchy generator ---------------------------

@pytest.mark.parametrize("k,p,lost", [(10, 4, [3, 12]), (10, 4, [0, 1, 10, 13]), (4, 2, [5]), (16, 4, [2, 7, 16, 19]),
                                      (32, 8, [0, 5, 33, 39, 31, 32, 34, 1]),
                                      (10, 4, [10]), (10, 4, [4]), (40, 2, [40])])   # one row, all ones: the XOR kernel
def test_repair_any_fragments_vs_oracle(rs, orc, torch, k, p, lost):
    L = 40_000 + 16
    data = splitmix64_bytes(900 + k, k * L).reshape(k, L)
    frags = np.concatenate([data, orc.matrix_mul(orc.gen_encoding_matrix(p, k), data)], axis=0)
    survivors = [i for i in range(k + p) if i not in lost][:k]
    with rs.RSContext(k, p) as ctx:
        ctx.set_repair(survivors, lost)
        d_s = to_dev(torch, frags[survivors])
        out = torch.empty((len(lost), L), dtype=torch.uint8, device="cuda")
        ctx.repair(d_s, L, out, L, L, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert (out.cpu().numpy() == frags[lost]).all()
        with pytest.raises(rs.RSGPUError):
            ctx.set_repair(survivors, [k + p])


def test_cauchy_generator_is_mds_and_roundtrips(rs, orc, torch):
    import itertools
    k, p, L = 10, 4, 4096 + 32
    data = splitmix64_bytes(1234, k * L).reshape(k, L)
    with rs.RSContext(k, p, generator="cauchy") as ctx:
        T = ctx.total_matrix()
        assert (T[:k] == np.eye(k, dtype=np.uint8)).all()
        for i in range(p):
            for j in range(k):
                assert orc.gf_mul(int(T[k + i, j]), (k + i) ^ j) == 1          # C[i][j] = 1 / ((k+i) + j)
        parity = gpu_encode(torch, ctx, data)
        assert (parity == orc.matrix_mul(T[k:], data)).all()
        frags = np.concatenate([data, parity], axis=0)
        # the sets that are singular under the reference's Vandermonde rows decode here,
        # and so does every other 4-erasure set (sampled) -- the code is MDS
        sets = [(0, 1, 2, 12), (0, 3, 4, 12), (4, 5, 9, 11)] + list(itertools.combinations(range(14), 4))[::37]
        for erased in sets:
            surv = [i for i in range(k + p) if i not in erased]
            ctx.set_survivors(surv)
            out = torch.empty((k, L), dtype=torch.uint8, device="cuda")
            ctx.decode(to_dev(torch, frags[surv]), L, out, L, L)
            torch.cuda.synchronize()
            assert (out.cpu().numpy() == data).all(), erased
    with pytest.raises(rs.RSGPUError):
        rs.RSContext(250, 10, generator="cauchy")          # k + p > 256


def test_contexts_on_concurrent_host_threads(rs, orc, torch):
    """include/rsgpu.h: one context per host thread.  Four threads, each with its own context, stream and code,
    run encode + erased decode concurrently on the same GPU (ctypes drops the GIL during the calls)."""
    import threading
    codes = [(10, 4, 300_007), (16, 4, 200_000), (6, 3, 150_016), (32, 8, 40_000)]
    errors = []

    def work(idx):
        try:
            k, p, L = codes[idx]
            data = splitmix64_bytes(31_000 + idx, k * L).reshape(k, L)
            want = orc.matrix_mul(orc.gen_encoding_matrix(p, k), data)
            st = torch.cuda.Stream()
            d = to_dev(torch, data)
            par = torch.empty((p, L), dtype=torch.uint8, device="cuda")
            lost = list(range(min(p, 3)))
            order = list(range(k, k + len(lost))) + list(range(len(lost), k))
            with rs.RSContext(k, p) as ctx:
                for it in range(15):
                    par.fill_(0xEE)
                    ctx.encode(d, L, par, L, L, st.cuda_stream)
                    st.synchronize()
                    assert (par.cpu().numpy() == want).all(), (idx, it)
                    ctx.set_survivors(order)
                    surv = torch.cat([par[:len(lost)], d[len(lost):]]).contiguous()
                    back = torch.empty((len(lost), L), dtype=torch.uint8, device="cuda")
                    ctx.decode_erased(surv, L, back, L, L, st.cuda_stream)
                    st.synchronize()
                    assert torch.equal(back, d[:len(lost)]), (idx, it)
        except Exception as e:          # noqa: BLE001 -- reported in the main thread
            errors.append((idx, repr(e)))

    ths = [threading.Thread(target=work, args=(i,)) for i in range(len(codes))]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    assert not errors, errors


# ---- randomized shapes (hypothesis): any k, p, length, pitch, misalignment, kernel variant -----------

def test_fuzz_encode_shapes_vs_oracle(rs, orc, torch):
    from hypothesis import given, settings, strategies as st, HealthCheck

    @settings(max_examples=60, deadline=None, derandomize=True, suppress_health_check=list(HealthCheck))
    @given(k=st.integers(1, 40), p=st.integers(1, 12), L=st.integers(1, 70_000), pad=st.sampled_from([0, 0, 16, 48, 5]),
           in_off=st.sampled_from([0, 0, 16, 3]), out_off=st.sampled_from([0, 0, 32, 7]),
           variant=st.sampled_from([0, 1, 5, 6, 7]), seed=st.integers(0, 2**31))
    def run(k, p, L, pad, in_off, out_off, variant, seed):
        data = np.random.default_rng(seed).integers(0, 256, (k, L), dtype=np.uint8)
        pitch = (L + 15) // 16 * 16 + pad
        with rs.RSContext(k, p) as ctx:
            ctx.set_tuning(variant=variant)
            got = gpu_encode(torch, ctx, data, pitch=pitch, out_pitch=pitch + 16, in_off=in_off, out_off=out_off)
        assert (got == orc.matrix_mul(orc.gen_encoding_matrix(p, k), data)).all(), (k, p, L, pad, in_off, out_off, variant)

    run()


def test_fuzz_unaligned_long_rows_vs_byte_kernel(rs, orc, torch):
    """Random codes, lengths just above the realigning kernels' threshold, random base offsets and pitch slack on both
    sides: the realigning kernels (byte tables or split-nibble tables, whichever the code gets) must write exactly what
    the byte kernel alone (variant 6) writes -- guard bytes included -- and a window of it is checked against the oracle."""
    from hypothesis import given, settings, strategies as st, HealthCheck

    @settings(max_examples=25, deadline=None, derandomize=True, suppress_health_check=list(HealthCheck))
    @given(k=st.integers(2, 48), p=st.sampled_from([1, 2, 3, 4, 5, 8, 9]), extra=st.integers(0, 5000), dpi=st.integers(0, 33),
           dpo=st.integers(0, 33), in_off=st.integers(0, 31), out_off=st.integers(0, 31), seed=st.integers(0, 2**31))
    def run(k, p, extra, dpi, dpo, in_off, out_off, seed):
        L = (1 << 20) + extra
        pi, po = L + dpi, L + dpo
        g = torch.Generator(device="cuda").manual_seed(seed)
        buf = torch.randint(0, 256, (k * pi + in_off + 64,), dtype=torch.uint8, device="cuda", generator=g)
        out = torch.full((p * po + out_off + 64,), 0xEE, dtype=torch.uint8, device="cuda")
        out6 = torch.full_like(out, 0xEE)
        stream = torch.cuda.current_stream().cuda_stream
        with rs.RSContext(k, p) as ctx:
            ctx.encode(buf.data_ptr() + in_off, pi, out.data_ptr() + out_off, po, L, stream)
            ctx.set_tuning(variant=6)
            ctx.encode(buf.data_ptr() + in_off, pi, out6.data_ptr() + out_off, po, L, stream)
            torch.cuda.synchronize()
        assert torch.equal(out, out6), (k, p, L, dpi, dpo, in_off, out_off)
        c0 = (seed % (L - 3000))
        data = buf[in_off:in_off + k * pi].view(k, pi)[:, c0:c0 + 3000].cpu().numpy()
        got = out[out_off:out_off + p * po].view(p, po)[:, c0:c0 + 3000].cpu().numpy()
        assert (got == orc.matrix_mul(orc.gen_encoding_matrix(p, k), data)).all(), (k, p, L, c0)

    run()


def test_fuzz_encode_batch_vs_per_stripe_calls(rs, orc, torch):
    """Random batch geometries: the one-call result must equal rsgpu_encode stripe by stripe (itself checked
    against the oracle above), and stripe 0 is checked against the oracle directly."""
    from hypothesis import given, settings, strategies as st, HealthCheck

    @settings(max_examples=40, deadline=None, derandomize=True, suppress_health_check=list(HealthCheck))
    @given(k=st.integers(1, 34), p=st.integers(1, 10), cols=st.integers(1, 3000), n=st.integers(1, 40),
           tail=st.sampled_from([0, 0, 0, 5]), layout=st.sampled_from(["stripe", "fragment"]),
           variant=st.sampled_from([0, 5, 5, 6]), seed=st.integers(0, 2**31))
    def run(k, p, cols, n, tail, layout, variant, seed):
        L = cols * 16 + tail
        stream = torch.cuda.current_stream().cuda_stream
        gen = torch.Generator(device="cuda").manual_seed(seed)
        din, dpitch, dsp = _batch_views(torch, k, k, L, n, layout, 0)
        din.random_(0, 256, generator=gen)
        out_a, ppitch, psp = _batch_views(torch, k, p, L, n, layout, 0xEE)
        out_b, _, _ = _batch_views(torch, k, p, L, n, layout, 0xEE)
        with rs.RSContext(k, p) as ctx:
            ctx.set_tuning(variant=variant)
            ctx.encode_batch(din.data_ptr(), dpitch, dsp, out_a.data_ptr(), ppitch, psp, L, n, stream)
            for s in range(n):
                ctx.encode(din.data_ptr() + s * dsp, dpitch, out_b.data_ptr() + s * psp, ppitch, L, stream)
            torch.cuda.synchronize()
        assert torch.equal(out_a, out_b), (k, p, L, n, layout, variant)    # untouched gaps (0xEE) included
        host_in, host_out = din.cpu().numpy(), out_a.cpu().numpy()
        d0 = np.stack([host_in[j * dpitch:j * dpitch + L] for j in range(k)])
        want = orc.matrix_mul(orc.gen_encoding_matrix(p, k), d0)
        for i in range(p):
            assert (host_out[i * ppitch:i * ppitch + L] == want[i]).all(), (k, p, L, n, layout, variant, i)

    run()


def test_batch_of_small_stripes_is_one_call(rs, orc, torch):
    """include/rsgpu.h: stripes stored fragment-major side by side are ONE encode of len = stripes * L."""
    k, p, L, n = 10, 4, 64 * 1024, 24
    stripes = [splitmix64_bytes(7000 + s, k * L).reshape(k, L) for s in range(n)]
    batch = np.concatenate(stripes, axis=1)                      # row j = fragment j of stripe 0, 1, ...
    with rs.RSContext(k, p) as ctx:
        got = gpu_encode(torch, ctx, batch)
    V = orc.gen_encoding_matrix(p, k)
    for s in (0, 7, n - 1):
        assert (got[:, s * L:(s + 1) * L] == orc.matrix_mul(V, stripes[s])).all()


# ---- per-row destinations, peer GPUs -----------------------------------------------------

BATCH_CASES = [
    # k, p, L, n_stripes, layout, variant
    (10, 4, 65_536, 37, "stripe", 0), (10, 4, 65_536, 37, "fragment", 0), (16, 4, 4_096 * 16, 64, "stripe", 0),
    (4, 2, 1 << 20, 5, "stripe", 0), (6, 3, 48_000, 19, "fragment", 0), (10, 8, 32_768, 21, "stripe", 0),
    (12, 4, 160, 1000, "stripe", 5), (9, 3, 20_000 * 16, 3, "stripe", 0), (20, 4, 16_384, 33, "stripe", 0),
    (10, 12, 16_384, 9, "stripe", 0),          # two row groups (8 + 4 rows): one batch launch each
    (7, 3, 65_536, 33, "stripe", 0), (13, 4, 32_768, 17, "fragment", 0), (32, 8, 16_384, 9, "stripe", 0),
    (17, 3, 65_536, 11, "stripe", 0), (5, 2, 65_536, 40, "fragment", 0), (23, 4, 16_384, 16, "stripe", 0),   # padded shapes
    (25, 3, 65_536, 11, "stripe", 0),          # no byte-table instance for k = 25: stripe-by-stripe fallback
    (10, 4, 10_007, 13, "stripe", 0),          # small ragged batch: stripe by stripe
    (10, 4, 10_007, 13, "stripe", 5), (16, 4, 65_541, 40, "fragment", 0), (6, 7, 4_111, 50, "stripe", 5),   # ragged len: + one tail launch
    (10, 4, 9, 100, "stripe", 5),              # shorter than one 16-byte column: stripe by stripe
    (10, 4, 65_536, 1, "stripe", 0), (10, 4, 256, 3, "stripe", 0),
]


def _batch_views(torch, k, rows, L, n, layout, fill):
    """device buffer + (row pitch, stripe pitch) for a batch of n stripes of `rows` rows of L bytes"""
    Lp = (L + 255) // 256 * 256
    if layout == "stripe":      # [stripe][row][Lp]
        pitch, spitch = Lp, rows * Lp + 512
    else:                       # [row][stripe][Lp]: rows of ALL stripes side by side
        pitch, spitch = n * Lp + 256, Lp
    size = (n - 1) * spitch + (rows - 1) * pitch + Lp if layout == "stripe" else rows * pitch
    buf = torch.full((size + 64,), fill, dtype=torch.uint8, device="cuda")
    return buf, pitch, spitch


@pytest.mark.parametrize("k,p,L,n,layout,variant", BATCH_CASES)
def test_encode_batch_vs_oracle(rs, orc, torch, k, p, L, n, layout, variant):
    data = splitmix64_bytes(9000 + k + n, n * k * L).reshape(n, k, L)
    din, dpitch, dsp = _batch_views(torch, k, k, L, n, layout, 0)
    dout, ppitch, psp = _batch_views(torch, k, p, L, n, layout, 0xEE)
    for s in range(n):
        for j in range(k):
            o = s * dsp + j * dpitch
            din[o:o + L] = to_dev(torch, data[s, j])
    with rs.RSContext(k, p) as ctx:
        ctx.set_tuning(variant=variant)
        n0 = ctx.launch_count()
        ctx.encode_batch(din.data_ptr(), dpitch, dsp, dout.data_ptr(), ppitch, psp, L, n, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        launches = ctx.launch_count() - n0
    V = orc.gen_encoding_matrix(p, k)
    host = dout.cpu().numpy()
    for s in range(n):
        want = orc.matrix_mul(V, data[s])
        for i in range(p):
            o = s * psp + i * ppitch
            assert (host[o:o + L] == want[i]).all(), f"stripe {s} parity {i}"
    if L >= 16 and n > 1 and k != 25 and (variant == 5 or L * n >= 2 << 20):
        per_group = 1 if L % 16 == 0 else 2        # + one byte-granular launch for the ragged tails of all stripes
        assert launches == per_group * ((p + 7) // 8), "an aligned batch is one launch per group of 8 parity rows"


def test_decode_erased_batch_roundtrip(rs, torch):
    k, p, L, n = 10, 4, 96 * 1024, 29
    stream = torch.cuda.current_stream().cuda_stream
    data = torch.randint(0, 256, (n, k, L), dtype=torch.uint8, device="cuda")
    parity = torch.empty((n, p, L), dtype=torch.uint8, device="cuda")
    order = [10, 1, 11, 3, 12, 5, 6, 13, 8, 9]            # natives 0, 2, 4, 7 lost
    with rs.RSContext(k, p) as ctx:
        ctx.encode_batch(data.data_ptr(), L, k * L, parity.data_ptr(), L, p * L, L, n, stream)
        frags = torch.cat([data, parity], dim=1)                          # (n, k+p, L)
        surv = frags[:, torch.tensor(order, device="cuda")].contiguous()  # (n, k, L)
        ctx.set_survivors(order)
        er = ctx.erased()
        assert er == [0, 2, 4, 7]
        out = torch.empty((n, len(er), L), dtype=torch.uint8, device="cuda")
        n0 = ctx.launch_count()
        ctx.decode_erased_batch(surv.data_ptr(), L, k * L, out.data_ptr(), L, len(er) * L, L, n, stream)
        torch.cuda.synchronize()
        assert ctx.launch_count() - n0 == 1
    assert torch.equal(out, data[:, torch.tensor(er, device="cuda")])


@pytest.mark.parametrize("k,p,L", [(10, 4, 100_000 + 16), (16, 4, 65_536 + 7), (7, 3, 33_333), (32, 8, 40_000), (5, 6, 4096)])
def test_encode_rows_same_device(rs, orc, torch, k, p, L):
    data = splitmix64_bytes(8100 + k, k * L).reshape(k, L)
    want = orc.matrix_mul(orc.gen_encoding_matrix(p, k), data)
    d = to_dev(torch, data)
    rows = [torch.full((L + 64,), 0x77, dtype=torch.uint8, device="cuda") for _ in range(p)]   # separate allocations
    with rs.RSContext(k, p) as ctx:
        ctx.encode_rows(d, L, rows, L, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
    for i in range(p):
        assert (rows[i][:L].cpu().numpy() == want[i]).all()
        assert bool((rows[i][L:] == 0x77).all())


def test_encode_rows_scatter_to_peer_gpu(rs, orc, torch):
    """Parity rows written straight into another GPU's memory over NVLink (needs >= 2 GPUs)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    k, p, L = 10, 4, 8 * 1024 * 1024 + 48
    data = splitmix64_bytes(8200, k * L).reshape(k, L)
    want = orc.matrix_mul(orc.gen_encoding_matrix(p, k), data)
    d = to_dev(torch, data)                                                      # on cuda:0
    rows = [torch.zeros(L, dtype=torch.uint8, device=f"cuda:{i % 2}") for i in range(p)]   # rows 1, 3 on the peer
    with rs.RSContext(k, p, device=0) as ctx:
        ctx.enable_peer(1)
        with torch.cuda.device(0):
            ctx.encode_rows(d, L, rows, L, torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
    for i in range(p):
        assert rows[i].device.index == i % 2
        assert (rows[i].cpu().numpy() == want[i]).all(), f"row {i}"
