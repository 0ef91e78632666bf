#!/usr/bin/env python3
"""Out-of-bounds probe (GPU box; run by tests/test_gpu_guard.py in a child process, because a fault kills the CUDA
context): every kernel family codes stripes whose input AND output buffers END EXACTLY AT AN UNMAPPED ADDRESS.

The buffers come from the CUDA virtual-memory API: a range of S + one granule of address space is reserved, only the
first S bytes are mapped, and the stripe is placed so that its last byte is byte S - 1.  One byte read or written past
the end of a buffer is then an illegal address, whatever a caching allocator would have hidden (torch rounds blocks up
and packs them into large segments; that is how a stale-pointer load in the cluster kernel's last partial warp survived
round 1).  Results are also compared with the same call on ordinary buffers.  Prints "guard ok: <n> cases".
"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
try:
    from cuda.bindings import driver as cu  # noqa: E402
except Exception as ex:      # no cuda-python in this environment: the caller skips
    print("guard probe unavailable:", ex)
    sys.exit(77)


def ck(res):
    err = res[0]
    if int(err) != 0:
        raise RuntimeError(f"CUDA driver error {err}")
    return res[1] if len(res) == 2 else res[1:]


class GuardedBuffer:
    """nbytes of device memory whose LAST byte is directly followed by unmapped address space."""

    def __init__(self, nbytes, device=0):
        prop = cu.CUmemAllocationProp()
        prop.type = cu.CUmemAllocationType.CU_MEM_ALLOCATION_TYPE_PINNED
        prop.location.type = cu.CUmemLocationType.CU_MEM_LOCATION_TYPE_DEVICE
        prop.location.id = device
        gran = ck(cu.cuMemGetAllocationGranularity(prop, cu.CUmemAllocationGranularity_flags.CU_MEM_ALLOC_GRANULARITY_MINIMUM))
        self.size = (max(nbytes, 1) + gran - 1) // gran * gran
        self.va = int(ck(cu.cuMemAddressReserve(self.size + gran, 0, 0, 0)))
        self.total = self.size + gran
        self.handle = ck(cu.cuMemCreate(self.size, prop, 0))
        ck(cu.cuMemMap(self.va, self.size, 0, self.handle, 0))
        acc = cu.CUmemAccessDesc()
        acc.location.type = cu.CUmemLocationType.CU_MEM_LOCATION_TYPE_DEVICE
        acc.location.id = device
        acc.flags = cu.CUmemAccess_flags.CU_MEM_ACCESS_FLAGS_PROT_READWRITE
        ck(cu.cuMemSetAccess(self.va, self.size, [acc], 1))
        self.ptr = self.va + self.size - nbytes          # the buffer ends where the mapping ends
        self.nbytes = nbytes

    def upload(self, t):
        assert t.numel() == self.nbytes and t.dtype == torch.uint8 and t.is_contiguous()
        ck(cu.cuMemcpyDtoD(self.ptr, t.data_ptr(), self.nbytes))

    def download(self):
        out = torch.empty(self.nbytes, dtype=torch.uint8, device="cuda")
        ck(cu.cuMemcpyDtoD(out.data_ptr(), self.ptr, self.nbytes))
        return out

    def free(self):
        ck(cu.cuMemUnmap(self.va, self.size))
        ck(cu.cuMemRelease(self.handle))
        ck(cu.cuMemAddressFree(self.va, self.total))


# k, p, L, pitch - L, variant
CASES = [
    (10, 4, (8 << 20) + 16 * 5, 0, 0),          # byte tables, column count not a multiple of the block / warp size
    (10, 4, (8 << 20) + 16 * 5 + 3, 0, 0),      # odd pitch: the realigning kernel + byte head / tail
    (10, 4, (1 << 20) + 1, 0, 0),
    (16, 4, (8 << 20) + 16 * 37, 0, 0),         # transposed bit-7 accumulators, L2 prefetch ahead of the end
    (15, 5, (8 << 20) + 16 * 3, 0, 0),          # padded shape, peeled XOR row
    (15, 5, (2 << 20) + 7, 0, 0),               # ... realigned
    (32, 8, (8 << 20) + 16 * 9, 0, 0),          # 64-bit entries
    (4, 2, (8 << 20) + 16, 0, 0),
    (24, 4, 16 * 777_777, 0, 8),                # cluster kernel, last warp partial
    (17, 3, 16 * 100_003, 0, 8),
    (32, 4, 16 * 200_001, 0, 8),
    (28, 4, 16 * 70_001 + 5, 11, 7),            # TMA-fed nibble kernel + byte tail
    (8, 4, (3 << 20) + 16 * 3, 0, 7),
    (27, 4, 16 * 65_539, 0, 6),                 # register-fed nibble kernel, padded batches
    (64, 4, 16 * 1536 * 3 + 16 * 77 + 9, 7, 9),  # multi-pass kernel + remainder + tail
    (40, 5, 16 * 1536 * 2, 0, 9),
    (28, 4, (1 << 20) + 3, 0, 0), (64, 8, (1 << 20) + 5, 0, 0), (33, 5, (1 << 20) + 9, 0, 0),   # realigning nibble kernel
    (10, 1, (1 << 20) + 16 * 3, 0, 0), (40, 1, (2 << 20) + 5, 11, 0),    # XOR kernel + tail
    (10, 4, 1000, 0, 0), (10, 4, 15, 0, 0),     # byte kernel only
]


def selftest(rs, st):
    """The guard must have teeth: a call that is TOLD a length 64 bytes longer than the buffers really are has to
    fault (exit code 3), otherwise a pass of the probe above would mean nothing."""
    k, p, L = 10, 4, (1 << 20)
    pitch = L + 64
    in_bytes, out_bytes = (k - 1) * pitch + L, (p - 1) * pitch + L
    gin, gout = GuardedBuffer(in_bytes), GuardedBuffer(out_bytes)
    gin.upload(torch.zeros(in_bytes, dtype=torch.uint8, device="cuda"))
    try:
        with rs.RSContext(k, p) as ctx:
            ctx.encode(gin.ptr, pitch, gout.ptr, pitch, L + 64, st)
            torch.cuda.synchronize()
    except Exception as ex:
        print("selftest: the overrun faulted as it must:", str(ex).splitlines()[0][:120])
        sys.exit(3)
    print("selftest: the overrun went UNNOTICED")
    sys.exit(0)


def main():
    rs = importlib.import_module("gpu-rscode_b200")
    rs.lib()
    torch.cuda.set_device(0)
    torch.zeros(1, device="cuda")           # primary context up before the driver API is used
    st = torch.cuda.current_stream().cuda_stream
    if "--selftest" in sys.argv:
        selftest(rs, st)
    n_ok = 0
    for k, p, L, dp, variant in CASES:
        pitch = L + dp
        # the last row has no padding behind it: the buffer ends with the last byte of the last row
        in_bytes, out_bytes = (k - 1) * pitch + L, (p - 1) * pitch + L
        g = torch.Generator(device="cuda").manual_seed(k * 1000 + p)
        flat = torch.randint(0, 256, (in_bytes,), dtype=torch.uint8, device="cuda", generator=g)
        want = torch.zeros(out_bytes, dtype=torch.uint8, device="cuda")
        gin, gout = GuardedBuffer(in_bytes), GuardedBuffer(out_bytes)
        gin.upload(flat)
        with rs.RSContext(k, p) as ctx:
            ctx.set_tuning(variant=variant)
            ctx.encode(flat.data_ptr(), pitch, want.data_ptr(), pitch, L, st)
            ctx.encode(gin.ptr, pitch, gout.ptr, pitch, L, st)
            torch.cuda.synchronize()        # an access past either buffer faults here
        got = gout.download()
        rows_w = [want[i * pitch:i * pitch + L] for i in range(p)]
        rows_g = [got[i * pitch:i * pitch + L] for i in range(p)]
        for i in range(p):
            assert torch.equal(rows_w[i], rows_g[i]), f"case {(k, p, L, dp, variant)} row {i} differs"
        gin.free()
        gout.free()
        n_ok += 1
        print(f"  ok  RS({k},{k + p}) L={L} pitch={pitch} variant={variant}", flush=True)
    print(f"guard ok: {n_ok} cases")


if __name__ == "__main__":
    main()
