"""rsgpu_encode_noalloc -- the entry point for callers that may not allocate GPU memory (an engine
living inside fixed device-memory arenas, or in an I/O path that cannot afford the device-wide stall
of cudaMalloc/cudaFree).  It keeps its tables in a fixed array inside the library instead of a context.

Checked here through the C ABI: its parity bytes must equal those of the ordinary context-based
rsgpu_encode, bit for bit (integer GF(2^8) work: tolerance zero), on two calls in a row that use
DIFFERENT code parameters, so the second call has to cope with what the first left in the shared
array.  Oversized codes must be refused without touching the output.

This file sorts last on purpose: with `pytest -x` every other test has run before it.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from gpu_lib import rs
    rs.lib()
    torch.cuda.set_device(0)
    return torch, rs


def _encode_both(torch, rs, k, p, L, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    data = torch.randint(0, 256, (k, L), dtype=torch.uint8, device="cuda", generator=g)
    st = torch.cuda.current_stream().cuda_stream
    want = torch.empty((p, L), dtype=torch.uint8, device="cuda")
    with rs.RSContext(k, p, device=0) as ctx:
        ctx.encode(data, L, want, L, L, st)
        torch.cuda.synchronize()
    got = torch.full((p, L), 0xA5, dtype=torch.uint8, device="cuda")
    rs.encode_noalloc(k, p, data, L, got, L, L, st)
    torch.cuda.synchronize()
    return got, want


def test_noalloc_matches_context_encode_two_calls_in_a_row(env):
    torch, rs = env
    # first call: wide code (64-bit table entries, 32 fragments) fills most of the shared array
    got, want = _encode_both(torch, rs, 32, 8, 1_000_003, seed=1)
    assert bool((got == want).all())
    # second call straight after, smaller code with other coefficients and a length that needs the byte tail
    got, want = _encode_both(torch, rs, 10, 4, 2_500_009, seed=2)
    assert bool((got == want).all())
    # and once more with the first shape: nothing stale from the second call either
    got, want = _encode_both(torch, rs, 32, 8, 4096, seed=3)
    assert bool((got == want).all())


def test_noalloc_reaches_the_fast_kernels_and_matches_the_oracle(env):
    """16-byte aligned stripes long enough for the kernels that expand the shared array's product tables into
    their own shared-memory tables (byte tables from 8 MiB per fragment, the TMA-fed kernel from 2 MiB), called
    back to back with different codes, so a stale table entry left by the previous call would show.  Compared
    with the context-based rsgpu_encode bit for bit and, on a window, with the CPU oracle directly."""
    torch, rs = env
    from oracle_lib import Oracle
    orc = Oracle()
    st = torch.cuda.current_stream().cuda_stream
    for k, p, L, seed in [(10, 4, (8 << 20) + 16, 11), (32, 8, 4 << 20, 12), (16, 4, (8 << 20) + 32, 13), (10, 4, 4 << 20, 14),
                          (12, 5, (8 << 20) + 3, 15)]:
        got, want = _encode_both(torch, rs, k, p, L, seed=seed)
        assert bool((got == want).all()), (k, p, L)
        g = torch.Generator(device="cuda").manual_seed(seed)
        data = torch.randint(0, 256, (k, L), dtype=torch.uint8, device="cuda", generator=g)     # the same bytes _encode_both drew
        for c0 in (0, L - 4096, (L // 2) & ~15):
            ref = orc.matrix_mul(orc.gen_encoding_matrix(p, k), data[:, c0:c0 + 4096].cpu().numpy())
            assert (got[:, c0:c0 + 4096].cpu().numpy() == ref).all(), (k, p, L, c0)


def test_noalloc_rejects_oversized_codes_and_bad_arguments(env):
    torch, rs = env
    d = torch.zeros((40, 256), dtype=torch.uint8, device="cuda")
    out = torch.full((16, 256), 0x3C, dtype=torch.uint8, device="cuda")
    for k, p in [(33, 4), (10, 9), (255, 1)]:
        with pytest.raises(rs.RSGPUError) as ei:
            rs.encode_noalloc(k, p, d, 256, out, 256, 256)
        assert ei.value.code == rs.RSGPU_ETOOBIG
        assert "k <= 32" in str(ei.value)
    for k, p in [(0, 2), (4, 0)]:
        with pytest.raises(rs.RSGPUError) as ei:
            rs.encode_noalloc(k, p, d, 256, out, 256, 256)
        assert ei.value.code == rs.RSGPU_EINVAL
    with pytest.raises(rs.RSGPUError):
        rs.encode_noalloc(4, 2, d, 128, out, 256, 256)      # pitch < len
    torch.cuda.synchronize()
    assert bool((out == 0x3C).all())                        # refused calls wrote nothing
