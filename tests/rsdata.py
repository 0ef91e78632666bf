"""Seeded synthetic data shared by tests, golden-vector generation and bench.py.

splitmix64(seed), 8 bytes little-endian per step.  First 16 bytes for
seed 1 are c15c0289ec2d0a9167ec8e65a18debbe.
"""
import numpy as np

_GAMMA = np.uint64(0x9E3779B97F4A7C15)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)


def splitmix64_bytes(seed: int, nbytes: int, first_step: int = 0) -> np.ndarray:
    """nbytes of the stream, starting at output number `first_step` (8 bytes per output)."""
    nwords = (nbytes + 7) // 8
    with np.errstate(over="ignore"):
        idx = np.arange(first_step + 1, first_step + nwords + 1, dtype=np.uint64)
        z = np.uint64(seed & 0xFFFFFFFFFFFFFFFF) + idx * _GAMMA
        z = (z ^ (z >> np.uint64(30))) * _M1
        z = (z ^ (z >> np.uint64(27))) * _M2
        z = z ^ (z >> np.uint64(31))
    return z.astype("<u8").view(np.uint8)[:nbytes].copy()


def pattern_bytes(mul: int, add: int, nbytes: int) -> np.ndarray:
    """file[i] = (mul*i + add) & 255 -- the tiny hand-checkable inputs of SURVEY section 8c."""
    i = np.arange(nbytes, dtype=np.int64)
    return ((mul * i + add) & 255).astype(np.uint8)
