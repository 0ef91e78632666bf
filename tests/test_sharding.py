"""CPU tests of the multi-GPU host logic: byte-range partition (C, host/sched.c), its Python mirror,
stripe dealing, and a world_size-2 gloo run in which each rank encodes only its own stripes (with the
oracle standing in for the kernel) and the results are combined exactly as bench.py combines them."""
import ctypes
import hashlib
import os
import socket

import numpy as np
import pytest

from gpu_lib import ROOT, rs

sharding = __import__("importlib").import_module("gpu-rscode_b200.sharding")


def _host_lib():
    if not os.path.exists(rs.HOST_LIB_PATH):
        rs.build()
    ctypes.CDLL(rs.LIB_PATH, mode=ctypes.RTLD_GLOBAL)
    H = ctypes.CDLL(rs.HOST_LIB_PATH)
    H.rs_partition.argtypes = [ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_size_t),
                               ctypes.POINTER(ctypes.c_size_t)]
    H.rs_partition.restype = None
    H.rs_plan_gpus.argtypes = [ctypes.c_size_t, ctypes.c_int]
    return H


@pytest.mark.parametrize("length", [0, 1, 255, 256, 4096, 10_000_019, (1 << 32) + 12345, 419_430_400])
@pytest.mark.parametrize("parts", [1, 2, 3, 4, 8])
def test_partition_covers_range_exactly(length, parts):
    H = _host_lib()
    pos = 0
    for g in range(parts):
        off, n = ctypes.c_size_t(), ctypes.c_size_t()
        H.rs_partition(length, parts, g, ctypes.byref(off), ctypes.byref(n))
        assert off.value == pos                      # contiguous, in order
        assert off.value % 256 == 0                  # every GPU keeps 16-byte aligned rows
        assert (off.value, n.value) == sharding.byte_range_of_rank(length, g, parts)
        pos += n.value
    assert pos == length


def test_plan_gpus_keeps_small_inputs_on_one_gpu():
    H = _host_lib()
    assert H.rs_plan_gpus(100, 8) == 1
    assert H.rs_plan_gpus(8 << 20, 8) == 2
    assert H.rs_plan_gpus(1 << 30, 8) == 8
    assert H.rs_plan_gpus(1 << 30, 0) == 1


def test_stripe_dealing_is_a_partition():
    for n, w in [(64, 1), (64, 2), (64, 8), (7, 4), (0, 2)]:
        seen = sorted(s for r in range(w) for s in sharding.stripes_of_rank(n, r, w))
        assert seen == list(range(n))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank, world, port, n_stripes, k, p, L, out_q):
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    from oracle_lib import Oracle
    from rsdata import splitmix64_bytes
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    orc = Oracle()
    V = orc.gen_encoding_matrix(p, k)
    digest = torch.zeros(n_stripes, 32, dtype=torch.uint8)
    for s in sharding.stripes_of_rank(n_stripes, rank, world):
        data = splitmix64_bytes(0xC4, k * L, first_step=s * k * L // 8).reshape(k, L)
        code = orc.matrix_mul(V, data)                       # the kernel's stand-in on CPU
        digest[s] = torch.frombuffer(bytearray(hashlib.sha256(code.tobytes()).digest()), dtype=torch.uint8)
    dist.barrier()
    # no data-path collective: only the bookkeeping is reduced (bench.py: max time over ranks)
    t = torch.tensor([10.0 + rank], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(digest, op=dist.ReduceOp.SUM)            # rows are disjoint, so SUM = union
    if rank == 0:
        out_q.put((float(t.item()), digest.numpy().tobytes()))
    dist.destroy_process_group()


def test_world2_gloo_stripe_sharding_matches_single_process():
    import torch.multiprocessing as mp
    from oracle_lib import Oracle
    from rsdata import splitmix64_bytes
    n_stripes, k, p, L, world = 6, 10, 4, 4096, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, world, port, n_stripes, k, p, L, q)) for r in range(world)]
    for pr in procs:
        pr.start()
    t_max, digests = q.get(timeout=120)
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    assert t_max == 11.0                                         # max over ranks, not rank 0's own
    orc = Oracle()
    V = orc.gen_encoding_matrix(p, k)
    want = b"".join(hashlib.sha256(orc.matrix_mul(
        V, splitmix64_bytes(0xC4, k * L, first_step=s * k * L // 8).reshape(k, L)).tobytes()).digest()
        for s in range(n_stripes))
    assert digests == want
    assert sharding.job_throughput_gbs(k * L, world, 3, 11.0) == pytest.approx(2 * k * L * 3 / 0.011 / 1e9)


def test_file_gpu_budget_and_visible_device_narrowing():
    """host/sched.c: one GPU per 32 GiB of file unless RS_GPUS says otherwise.  The RS command line (and only it:
    encode_file / decode_file leave the environment of an embedding program alone) narrows CUDA_VISIBLE_DEVICES
    to the first n devices (of the user's list if any) before CUDA starts.  Run in a child: it edits the environment."""
    import subprocess
    import sys
    code = r'''
import ctypes, os, sys
os.environ.pop("CUDA_VISIBLE_DEVICES", None); os.environ.pop("RS_GPUS", None)
ctypes.CDLL(sys.argv[1], mode=ctypes.RTLD_GLOBAL)
H = ctypes.CDLL(sys.argv[2])
H.rs_file_gpu_budget.argtypes = [ctypes.c_longlong]
libc = ctypes.CDLL(None); libc.getenv.restype = ctypes.c_char_p
assert H.rs_file_gpu_budget(0) == 1 and H.rs_file_gpu_budget(64 << 20) == 1
assert H.rs_file_gpu_budget(32 << 30) == 1 and H.rs_file_gpu_budget((32 << 30) + 1) == 2
assert H.rs_file_gpu_budget(1 << 40) == 32
libc.setenv(b"RS_GPUS", b"8", 1)
assert H.rs_file_gpu_budget(10) == 8
import tempfile
d = tempfile.mkdtemp()
f = os.path.join(d, "file.bin").encode(); open(f, "wb").write(b"x" * 1000)
open(f + b".METADATA", "w").write("1000\n2 4\n")
H.rs_cli_prepare_gpus.argtypes = [ctypes.c_char_p, ctypes.c_int]
libc.setenv(b"RS_GPUS", b"3", 1)
H.rs_cli_prepare_gpus(f, 0); assert libc.getenv(b"CUDA_VISIBLE_DEVICES") == b"0,1,2"
libc.setenv(b"CUDA_VISIBLE_DEVICES", b"4,5,6,7", 1); libc.setenv(b"RS_GPUS", b"2", 1); H.rs_cli_prepare_gpus(f, 1)
assert libc.getenv(b"CUDA_VISIBLE_DEVICES") == b"4,5"
libc.setenv(b"CUDA_VISIBLE_DEVICES", b"6", 1); libc.setenv(b"RS_GPUS", b"4", 1); H.rs_cli_prepare_gpus(f, 0)
assert libc.getenv(b"CUDA_VISIBLE_DEVICES") == b"6"
libc.unsetenv(b"RS_GPUS"); libc.setenv(b"CUDA_VISIBLE_DEVICES", b"0,1,2,3", 1); H.rs_cli_prepare_gpus(f, 0)
assert libc.getenv(b"CUDA_VISIBLE_DEVICES") == b"0"                     # 1000 bytes: one GPU
H.rs_cli_prepare_gpus(os.path.join(d, "missing").encode(), 0)          # unknown size: environment untouched
assert libc.getenv(b"CUDA_VISIBLE_DEVICES") == b"0"
print("ok")
'''
    r = subprocess.run([sys.executable, "-c", code, rs.LIB_PATH, rs.HOST_LIB_PATH], stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stdout
