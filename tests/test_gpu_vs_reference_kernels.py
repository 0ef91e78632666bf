"""Second checker on the GPU box: the reference's OWN CUDA kernels (src/matrix.cu, unmodified, recompiled
for sm_100a into oracle/_ref/libgpurs_ref.so by oracle/Makefile) against this engine's kernels on the
same device buffers.  Bit-exact.  Skipped when that library was not built (it needs the reference tree
and nvcc at build time; the prebuilt file travels to the GPU box like the other oracle/_ref outputs)."""
import numpy as np
import pytest

from oracle_lib import Oracle, ReferenceGPU, have_reference_gpu
from rsdata import splitmix64_bytes

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not have_reference_gpu(), reason="oracle/_ref/libgpurs_ref.so not built")]


@pytest.fixture(scope="module")
def torch():
    import torch as t
    if not t.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return t


@pytest.fixture(scope="module")
def rs(torch):
    from gpu_lib import rs as m
    m.lib()
    return m


@pytest.fixture(scope="module")
def ref(torch):
    torch.zeros(1, device="cuda")          # the runtime context both libraries share
    return ReferenceGPU()


def dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


# (k, p, L): word kernel when L % 8 == 0 (matrix.cu:796), byte kernel otherwise (matrix.cu:811).
# Every L here is a whole number of the reference's tiles (128/p columns of 8 bytes, or of 1 byte in the
# byte kernel): with a partial last tile its kernels call __syncthreads() inside `if (row < n && col < m)`
# (matrix.cu:273-316), which deadlocks on sm_70 and later -- observed on B200 with RS(10,14), L = 12296:
# the launch never returns.  This engine has no such restriction (tests/test_gpu_parity.py covers ragged L).
@pytest.mark.parametrize("k,p,L", [(4, 2, 1 << 20), (10, 4, 1 << 20), (10, 4, 3 * 4096 + 256), (16, 4, 1 << 18),
                                     (32, 8, 1 << 16), (6, 3, 42 * 8 * 15), (10, 3, 42 * 505), (4, 5, 25 * 1001)])
def test_encode_same_bytes_as_reference_kernels(rs, ref, torch, k, p, L):
    data = dev(torch, splitmix64_bytes(4242 + k + p, k * L).reshape(k, L))
    with rs.RSContext(k, p) as ctx:
        coeff = dev(torch, ctx.total_matrix()[k:])
        ours = torch.full((p, L), 0xEE, dtype=torch.uint8, device="cuda")
        ctx.encode(data.data_ptr(), L, ours.data_ptr(), L, L, torch.cuda.current_stream().cuda_stream)
    theirs = torch.full((p, L), 0x11, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    ms = ref.encode_chunk(data.data_ptr(), coeff.data_ptr(), theirs.data_ptr(), k, p, L)
    torch.cuda.synchronize()
    assert ms > 0
    assert torch.equal(ours, theirs)


@pytest.mark.parametrize("k,p,order,L", [(4, 2, [0, 1, 4, 5], 1 << 18), (10, 4, [10, 11, 12, 13, 4, 5, 6, 7, 8, 9], 1 << 18),
                                           (16, 4, [19, 18] + list(range(2, 16)), 1 << 16), (3, 2, [3, 1, 2], 42 * 953)])
def test_decode_same_bytes_as_reference_kernels(rs, ref, torch, k, p, order, L):
    """The reference's decode_chunk multiplies by whatever inverse it is handed (src/decode.cu:166); given
    this engine's inverse both must return the original natives."""
    natives = splitmix64_bytes(999 + k, k * L).reshape(k, L)
    data = dev(torch, natives)
    stream = torch.cuda.current_stream().cuda_stream
    with rs.RSContext(k, p) as ctx:
        parity = torch.empty((p, L), dtype=torch.uint8, device="cuda")
        ctx.encode(data.data_ptr(), L, parity.data_ptr(), L, L, stream)
        frags = torch.cat([data, parity])
        surv = frags[torch.tensor(order, device="cuda")].contiguous()
        ctx.set_survivors(order)
        inv = dev(torch, ctx.decode_matrix())
        ours = torch.empty((k, L), dtype=torch.uint8, device="cuda")
        ctx.decode(surv.data_ptr(), L, ours.data_ptr(), L, L, stream)
    theirs = torch.full((k, L), 0x11, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    ref.decode_chunk(theirs.data_ptr(), inv.data_ptr(), surv.data_ptr(), k, p, L)
    torch.cuda.synchronize()
    assert torch.equal(ours, theirs)
    assert torch.equal(ours, data)


# the reference kernel has no bounds check (matrix.cu:752-759): only shapes its launch covers exactly
@pytest.mark.parametrize("k,p", [(4, 2), (10, 4), (16, 4), (32, 8), (16, 16)])
def test_generator_rows_same_as_reference_kernel(rs, ref, torch, k, p):
    theirs = torch.zeros(p * k, dtype=torch.uint8, device="cuda")
    ref.gen_encoding_matrix(theirs.data_ptr(), p, k)
    with rs.RSContext(k, p) as ctx:
        ours = ctx.total_matrix()[k:]
    assert (theirs.cpu().numpy().reshape(p, k) == ours).all()
    assert (ours == Oracle().gen_encoding_matrix(p, k)).all()
