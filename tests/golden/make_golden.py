#!/usr/bin/env python3
"""Generate tests/golden/golden.json by RUNNING THE REFERENCE ITSELF (src/cpu-rs.c, unmodified).

Run in the build container (needs /root/reference):
    make -C oracle && python tests/golden/make_golden.py

The reference is used two ways: oracle/_ref/libcpurs_ref.so (cpu-rs.c with -Dmain=cpu_rs_main,
called through ctypes) for tables / matrices / matrix_mul / invert_matrix, and the stand-alone
oracle/_ref/CPU-RS binary for the file-level encode / decode (fragments, METADATA, decoded file).
Nothing here uses our own oracle or engine, so the fixtures pin both.  The file is small (hex for
tiny cases, sha256 for the rest) and is committed; the GPU box never needs /root/reference.
"""
import hashlib
import itertools
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_lib import REF_BIN, Reference  # noqa: E402
from rsdata import pattern_bytes, splitmix64_bytes  # noqa: E402


def sha(b) -> str:
    return hashlib.sha256(bytes(b)).hexdigest()


def make_input(spec) -> np.ndarray:
    if spec["gen"] == "pattern":
        return pattern_bytes(spec["mul"], spec["add"], spec["nbytes"])
    return splitmix64_bytes(spec["seed"], spec["nbytes"])


ENCODE_CASES = [
    # tiny, hex-checkable
    dict(k=4, p=2, gen="pattern", mul=1, add=0, nbytes=32),
    dict(k=10, p=4, gen="pattern", mul=7, add=3, nbytes=80),
    dict(k=4, p=2, gen="pattern", mul=31, add=5, nbytes=13),     # padding of the last native
    dict(k=3, p=2, gen="pattern", mul=5, add=1, nbytes=1),       # 1-byte file, two all-pad natives
    dict(k=5, p=3, gen="seed", seed=11, nbytes=4097),
    dict(k=10, p=4, gen="seed", seed=3, nbytes=100_003),
    dict(k=16, p=4, gen="seed", seed=4, nbytes=262_144 + 5),
    dict(k=32, p=8, gen="seed", seed=5, nbytes=524_288 + 17),
    dict(k=6, p=6, gen="seed", seed=6, nbytes=70_001),
    dict(k=20, p=12, gen="seed", seed=8, nbytes=50_021),
    dict(k=10, p=4, gen="seed", seed=1, nbytes=10_485_773),      # SURVEY 8c seeded fixture
    dict(k=4, p=2, gen="seed", seed=2, nbytes=4 * 1024 * 1024),  # SURVEY 8c seeded fixture
]

# (encode-case index, conf order) ; orders chosen inside and outside the reference-correct domain
DECODE_CASES = [
    (2, [0, 1, 2, 3]),
    (2, [0, 1, 4, 5]),
    (2, [4, 5, 2, 3]),
    (2, [0, 1, 5, 3]),
    (2, [0, 1, 3, 5]),          # reference FAILS here (zero on the running diagonal)
    (2, [2, 3, 4, 5]),          # the unit-test.sh pattern: reference FAILS
    (5, [10, 11, 12, 13, 4, 5, 6, 7, 8, 9]),
    (5, [0, 11, 2, 3, 4, 13, 6, 7, 8, 9]),
    (6, [16, 17, 18, 19] + list(range(4, 16))),
    (7, list(range(32, 40)) + list(range(8, 32))),
    (4, [0, 5, 2, 7, 4]),
]

INVERSE_CASES = [
    (4, 2, [0, 1, 4, 5]), (4, 2, [4, 5, 2, 3]), (4, 2, [0, 1, 5, 3]),
    (10, 4, [10, 11, 12, 13, 4, 5, 6, 7, 8, 9]),
    (10, 4, [0, 1, 2, 10, 4, 5, 11, 7, 8, 13]),
    (16, 4, [16, 17, 18, 19] + list(range(4, 16))),
    (16, 4, list(range(0, 12)) + [16, 17, 18, 19]),
    (32, 8, list(range(32, 40)) + list(range(8, 32))),
    (32, 8, list(range(0, 24)) + list(range(32, 40))),
]


def main():
    ref = Reference()
    out = {"generator": "tests/golden/make_golden.py", "source": "reference src/cpu-rs.c via oracle/_ref"}

    out["log"] = ref.log_table().tobytes().hex()
    out["alog"] = ref.alog_table().tobytes().hex()
    mul = np.array([[ref.gf_mul(a, b) for b in range(256)] for a in range(256)], np.uint8)
    div = np.array([[ref.gf_div(a, b) for b in range(256)] for a in range(256)], np.uint8)
    pw = np.array([[ref.gf_pow(a, b) for b in range(256)] for a in range(256)], np.uint8)
    out["gf_mul_sha256"] = sha(mul.tobytes())
    out["gf_div_sha256"] = sha(div.tobytes())
    out["gf_pow_sha256"] = sha(pw.tobytes())
    out["gf_mul_rows"] = {str(a): mul[a].tobytes().hex() for a in (2, 3, 29, 142, 255)}

    out["vandermonde"] = {}
    for k, p in [(4, 2), (10, 4), (16, 4), (32, 8), (5, 3), (20, 12), (64, 2), (128, 3), (255, 1)]:
        out["vandermonde"][f"{k},{p}"] = ref.gen_encoding_matrix(p, k).tobytes().hex()

    # matrix_mul on raw buffers (no files): C = V * data for odd lengths
    out["matmul"] = []
    for k, p, m, seed in [(4, 2, 1, 21), (10, 4, 15, 22), (10, 4, 1000, 23), (16, 4, 4099, 24),
                          (32, 8, 2051, 25), (7, 9, 333, 26)]:
        data = splitmix64_bytes(seed, k * m).reshape(k, m)
        V = ref.gen_encoding_matrix(p, k)
        code = ref.matrix_mul(V, data)
        out["matmul"].append(dict(k=k, p=p, m=m, seed=seed, code_sha256=sha(code.tobytes()),
                                  code_head=code[:, :8].tobytes().hex() if m >= 8 else code.tobytes().hex()))

    # inverse of survivor sub-matrices through the reference's invert_matrix
    out["inverse"] = []
    for k, p, order in INVERSE_CASES:
        T = ref.gen_total_encoding_matrix(k, p)
        sub = T[order]
        inv = ref.invert_matrix(sub)
        ident = ref.matrix_mul(sub, inv)
        out["inverse"].append(dict(k=k, p=p, order=order, inverse=inv.tobytes().hex(),
                                   is_true_inverse=bool((ident == np.eye(k, dtype=np.uint8)).all())))

    # exhaustive RS(4,6): every ordered choice is too many; all 15 survivor sets in "natural" order
    nat = []
    for k, p in [(4, 2)]:
        T = ref.gen_total_encoding_matrix(k, p)
        for erased in itertools.combinations(range(k), 2):
            order = list(range(k))
            for slot, par in zip(erased, range(k, k + p)):
                order[slot] = par
            inv = ref.invert_matrix(T[order])
            nat.append(dict(k=k, p=p, order=order, inverse=inv.tobytes().hex()))
    out["inverse_natural_rs46"] = nat

    # file level through the CPU-RS binary
    out["encode"] = []
    enc_dirs = []
    tmp = tempfile.mkdtemp(prefix="golden_")
    for ci, spec in enumerate(ENCODE_CASES):
        d = os.path.join(tmp, f"c{ci}")
        os.makedirs(d)
        data = make_input(spec)
        with open(os.path.join(d, "f.bin"), "wb") as f:
            f.write(data.tobytes())
        k, p = spec["k"], spec["p"]
        subprocess.run([REF_BIN, "-k", str(k), "-n", str(k + p), "-e", "f.bin"], cwd=d, check=True,
                       stdout=subprocess.DEVNULL)
        frags = [open(os.path.join(d, f"_{i}_f.bin"), "rb").read() for i in range(k + p)]
        rec = dict(spec)
        rec["input_sha256"] = sha(data.tobytes())
        rec["chunk"] = len(frags[0])
        rec["metadata"] = open(os.path.join(d, "f.bin.METADATA")).read()
        rec["fragment_sha256"] = [sha(fr) for fr in frags]
        if len(frags[0]) <= 16:
            rec["fragment_hex"] = [fr.hex() for fr in frags]
        out["encode"].append(rec)
        enc_dirs.append(d)

    out["decode"] = []
    for ci, order in DECODE_CASES:
        d = enc_dirs[ci]
        spec = ENCODE_CASES[ci]
        with open(os.path.join(d, "conf"), "w") as f:
            for idx in order:
                f.write(f"_{idx}_f.bin\n")
        subprocess.run([REF_BIN, "-d", "-i", "f.bin", "-c", "conf", "-o", "out.bin"], cwd=d, check=True,
                       stdout=subprocess.DEVNULL)
        got = open(os.path.join(d, "out.bin"), "rb").read()
        orig = open(os.path.join(d, "f.bin"), "rb").read()
        out["decode"].append(dict(encode_case=ci, k=spec["k"], p=spec["p"], order=order,
                                  output_sha256=sha(got), reference_correct=bool(got == orig)))

    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", os.path.join(HERE, "golden.json"), os.path.getsize(os.path.join(HERE, "golden.json")), "bytes")


if __name__ == "__main__":
    main()
