"""Stripe sharding across ranks / GPUs -- host-side logic only (no CUDA here).

The coding path needs no exchange between GPUs: whole stripes are dealt round-robin
to ranks (device-resident batches, bench.py) and a single stripe is cut into contiguous byte ranges of
every fragment (files, host/sched.c: rs_partition -- mirrored here so both can be tested together).
"""


def stripes_of_rank(n_stripes: int, rank: int, world: int):
    """Rank g owns stripes s with s % world == g."""
    return list(range(rank, n_stripes, world))


def byte_range_of_rank(length: int, rank: int, world: int):
    """(offset, size) of the byte range of every fragment that rank owns; boundaries on 256 bytes,
    last rank takes the remainder.  Same arithmetic as rs_partition() in host/sched.c."""
    per = (length // world) // 256 * 256
    off = rank * per
    return off, (length - off if rank == world - 1 else per)


def job_throughput_gbs(units_per_rank_bytes: int, world: int, steps: int, max_rank_ms: float) -> float:
    """Whole-job GB/s: bytes all ranks processed divided by the slowest rank's time."""
    return world * units_per_rank_bytes * steps / (max_rank_ms / 1e3) / 1e9
