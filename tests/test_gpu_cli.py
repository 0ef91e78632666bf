"""Drop-in tests of the C host layer (RS binary = include/encode.h + include/decode.h + CLI) on a GPU:
same fragment files, METADATA and conf handling as the reference's programs, checked against the
reference's own CPU-RS binary (oracle/_ref, built from its unmodified cpu-rs.c) when that is present and
against our C oracle otherwise."""
import json
import os
import subprocess

import numpy as np
import pytest

from oracle_lib import REF_BIN, Oracle, have_reference
from rsdata import pattern_bytes, splitmix64_bytes

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "golden.json")))


@pytest.fixture(scope="module")
def rs_bin():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from gpu_lib import rs
    assert os.path.exists(rs.RS_BIN), "gpu-rscode_b200/RS not built"
    return rs.RS_BIN


@pytest.fixture(scope="module")
def orc():
    return Oracle()


def run(cmd, cwd, ok=(0,)):
    r = subprocess.run(cmd, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode in ok, r.stdout
    return r


def cpu_encode(orc, d, fname, k, p):
    """Reference CPU encode of d/fname: the real CPU-RS binary if available, else our oracle port."""
    if have_reference():
        run([REF_BIN, "-k", str(k), "-n", str(k + p), "-e", fname], d)
    else:
        cwd = os.getcwd()
        os.chdir(d)
        try:
            assert orc.encode_file(fname, k, p) == 0
        finally:
            os.chdir(cwd)


def write_conf(d, fname, order, conf="conf"):
    with open(os.path.join(d, conf), "w") as f:
        for i in order:
            f.write(f"_{i}_{fname}\n")


CASES = [(4, 2, 13), (4, 2, 32), (10, 4, 80), (3, 2, 1), (5, 3, 4097), (10, 4, 1_000_003), (16, 4, 262_149), (32, 8, 524_305),
         (4, 2, 9_000_000)]


@pytest.mark.parametrize("k,p,size", CASES)
def test_cli_encode_matches_cpu_reference(rs_bin, orc, tmp_path, k, p, size):
    data = splitmix64_bytes(size + k, size)
    a, b = tmp_path / "gpu", tmp_path / "cpu"
    a.mkdir(), b.mkdir()
    (a / "f.bin").write_bytes(data.tobytes())
    (b / "f.bin").write_bytes(data.tobytes())
    out = run([rs_bin, "-k", str(k), "-n", str(k + p), "-e", "f.bin"], a).stdout
    assert "Total GPU encoding time using multiple devices:" in out
    cpu_encode(orc, str(b), "f.bin", k, p)
    for i in range(k + p):
        assert (a / f"_{i}_f.bin").read_bytes() == (b / f"_{i}_f.bin").read_bytes(), f"fragment {i}"
    # METADATA: first two lines as the CPU program writes them, then the GPU writer's matrix rows
    meta = (a / "f.bin.METADATA").read_text()
    cpu_meta = (b / "f.bin.METADATA").read_text()
    assert meta.startswith(cpu_meta)
    T = orc.gen_total_encoding_matrix(k, p)
    want = cpu_meta + "".join("".join(f"{int(v)} " for v in row) + "\n" for row in T)   # src/encode.cu:70-98
    assert meta == want


def test_cli_encode_tiny_golden_hex(rs_bin, tmp_path):
    spec = GOLD["encode"][2]       # RS(4,6), 13 bytes, padded last native
    (tmp_path / "f.bin").write_bytes(pattern_bytes(spec["mul"], spec["add"], spec["nbytes"]).tobytes())
    run([rs_bin, "-k", "4", "-n", "6", "-e", "f.bin"], tmp_path)
    assert [(tmp_path / f"_{i}_f.bin").read_bytes().hex() for i in range(6)] == spec["fragment_hex"]


DEC = [
    (4, 2, 100_003, [0, 1, 4, 5]),
    (4, 2, 100_003, [2, 3, 4, 5]),        # unit-test.sh order: the reference writes wrong data here
    (4, 2, 100_003, [0, 1, 3, 5]),        # ditto
    (4, 2, 100_003, [5, 4, 1, 0]),
    (10, 4, 1_000_003, [10, 11, 12, 13, 4, 5, 6, 7, 8, 9]),
    (10, 4, 77, list(range(10))),
    (16, 4, 300_001, [16, 17, 18, 19] + list(range(4, 16))),
    (32, 8, 640_007, list(range(32, 40)) + list(range(8, 32))),
    (3, 2, 1, [3, 4, 2]),
]


@pytest.mark.parametrize("k,p,size,order", DEC, ids=lambda v: str(v)[:20])
def test_cli_roundtrip_and_cross_decode(rs_bin, orc, tmp_path, k, p, size, order):
    data = splitmix64_bytes(99 + size, size)
    a, b = tmp_path / "gpu", tmp_path / "cpu"
    a.mkdir(), b.mkdir()
    (a / "f.bin").write_bytes(data.tobytes())
    (b / "f.bin").write_bytes(data.tobytes())
    run([rs_bin, "-k", str(k), "-n", str(k + p), "-e", "f.bin"], a)
    cpu_encode(orc, str(b), "f.bin", k, p)
    write_conf(a, "f.bin", order)
    write_conf(b, "f.bin", order)
    # our decoder on our fragments (GPU-writer METADATA with matrix)
    out = run([rs_bin, "-d", "-i", "f.bin", "-c", "conf", "-o", "out.bin"], a).stdout
    assert "Total GPU decoding time using multiple devices:" in out
    assert (a / "out.bin").read_bytes() == data.tobytes()
    # our decoder on the CPU program's fragments (two-line METADATA, matrix regenerated)
    run([rs_bin, "-d", "-i", "f.bin", "-c", "conf", "-o", "out.bin"], b)
    assert (b / "out.bin").read_bytes() == data.tobytes()
    # the reference CPU decoder on OUR fragments + METADATA, wherever the reference itself is right
    sub = orc.gen_total_encoding_matrix(k, p)[order]
    _, st = orc.invert_matrix(sub)
    if st == 0:
        if have_reference():
            run([REF_BIN, "-d", "-i", "f.bin", "-c", "conf", "-o", "ref_out.bin"], a)
        else:
            cwd = os.getcwd()
            os.chdir(a)
            try:
                assert orc.decode_file("f.bin", "conf", "ref_out.bin")[0] == 0
            finally:
                os.chdir(cwd)
        assert (a / "ref_out.bin").read_bytes() == data.tobytes()


def test_cli_decode_overwrites_original_without_o(rs_bin, tmp_path):
    data = splitmix64_bytes(5, 50_000)
    (tmp_path / "f.bin").write_bytes(data.tobytes())
    run([rs_bin, "-k", "4", "-n", "6", "-e", "f.bin"], tmp_path)
    (tmp_path / "f.bin").write_bytes(b"garbage")
    write_conf(tmp_path, "f.bin", [4, 1, 5, 3])
    run([rs_bin, "-d", "-i", "f.bin", "-c", "conf"], tmp_path)          # src/decode.cu:410-417
    assert (tmp_path / "f.bin").read_bytes() == data.tobytes()


def test_cli_singular_survivors_reported(rs_bin, tmp_path):
    data = splitmix64_bytes(6, 10_000)
    (tmp_path / "f.bin").write_bytes(data.tobytes())
    run([rs_bin, "-k", "10", "-n", "14", "-e", "f.bin"], tmp_path)
    surv = [i for i in range(14) if i not in (0, 1, 2, 12)]            # singular set
    write_conf(tmp_path, "f.bin", surv)
    r = run([rs_bin, "-d", "-i", "f.bin", "-c", "conf", "-o", "out.bin"], tmp_path, ok=(1,))
    assert "Cannot decode" in r.stdout
    assert not (tmp_path / "out.bin").exists()


def test_cli_flag_rules_and_errors(rs_bin, tmp_path):
    r = run([rs_bin, "-i", "x", "-d"], tmp_path)                       # -i before -d -> usage, exit 0 (src/main.c:104)
    assert r.stdout.startswith("Usage:")
    assert run([rs_bin, "-h"], tmp_path).stdout.startswith("Usage:")
    r = run([rs_bin, "-k", "4", "-n", "6", "-e", "missing.bin"], tmp_path)
    assert "Cannot open source file!" in r.stdout                      # src/encode.cu:308-312, exit(0)
    r = run([rs_bin, "-d", "-i", "missing.bin", "-c", "conf"], tmp_path)
    assert "Cannot open metadata file!" in r.stdout
    (tmp_path / "g.bin").write_bytes(b"0123456789")
    r = run([rs_bin, "-K", "2", "-N", "3", "-E", "g.bin", "-p", "64", "-s", "2"], tmp_path)   # upper-case flags, -p/-s
    assert "Valid grid size" not in r.stdout
    assert (tmp_path / "_2_g.bin").read_bytes() == bytes(a ^ b for a, b in zip(b"01234", b"56789"))
    r = run([rs_bin, "-k", "2", "-n", "3", "-e", "g.bin"], tmp_path)
    assert "Valid grid size: (0, 65535]" in r.stdout                   # printed whenever -p is absent, src/encode.cu:351-355


@pytest.mark.parametrize("di", range(len(GOLD["decode"])))
def test_cli_decode_golden(rs_bin, tmp_path, di):
    case = GOLD["decode"][di]
    spec = GOLD["encode"][case["encode_case"]]
    raw = pattern_bytes(spec["mul"], spec["add"], spec["nbytes"]) if spec["gen"] == "pattern" \
        else splitmix64_bytes(spec["seed"], spec["nbytes"])
    (tmp_path / "f.bin").write_bytes(raw.tobytes())
    run([rs_bin, "-k", str(spec["k"]), "-n", str(spec["k"] + spec["p"]), "-e", "f.bin"], tmp_path)
    write_conf(tmp_path, "f.bin", case["order"])
    run([rs_bin, "-d", "-i", "f.bin", "-c", "conf", "-o", "out.bin"], tmp_path)
    got = (tmp_path / "out.bin").read_bytes()
    assert got == raw.tobytes()                    # always the true data ...
    if case["reference_correct"]:                  # ... which is what the reference wrote wherever it is right
        import hashlib
        assert hashlib.sha256(got).hexdigest() == case["output_sha256"]


def test_cli_cauchy_generator_decodes_reference_singular_set(rs_bin, tmp_path):
    """RS_GENERATOR=cauchy: the matrix travels in METADATA; a survivor set that is singular under the
    reference's Vandermonde rows decodes."""
    data = splitmix64_bytes(8, 123_457)
    (tmp_path / "f.bin").write_bytes(data.tobytes())
    env = dict(os.environ, RS_GENERATOR="cauchy")
    r = subprocess.run([rs_bin, "-k", "10", "-n", "14", "-e", "f.bin"], cwd=tmp_path, env=env, stdout=subprocess.PIPE, text=True)
    assert r.returncode == 0
    rows = (tmp_path / "f.bin.METADATA").read_text().splitlines()
    assert rows[1] == "4 10" and len(rows) == 2 + 14 and rows[12].split()[0] != "1"      # not the all-ones Vandermonde row
    write_conf(tmp_path, "f.bin", [i for i in range(14) if i not in (0, 1, 2, 12)])
    run([rs_bin, "-d", "-i", "f.bin", "-c", "conf", "-o", "out.bin"], tmp_path)
    assert (tmp_path / "out.bin").read_bytes() == data.tobytes()


def test_cli_streaming_many_slabs_ring_reuse(rs_bin, tmp_path):
    """A file big enough that the slab ring (GPUs + 2 slots) is reused many times: RS(32,40), 230 MB ->
    chunk 7.2 MB, 2 MiB slabs."""
    import hashlib
    size = 230_000_003
    data = splitmix64_bytes(0xC5, size)
    (tmp_path / "big.bin").write_bytes(data.tobytes())
    want = hashlib.sha256(data.tobytes()).hexdigest()
    del data
    run([rs_bin, "-k", "32", "-n", "40", "-e", "big.bin"], tmp_path)
    assert (tmp_path / "_39_big.bin").stat().st_size == (size + 31) // 32
    order = list(range(32, 40)) + list(range(8, 32))
    write_conf(tmp_path, "big.bin", order)
    run([rs_bin, "-d", "-i", "big.bin", "-c", "conf", "-o", "out.bin"], tmp_path)
    h = hashlib.sha256()
    with open(tmp_path / "out.bin", "rb") as f:
        for blk in iter(lambda: f.read(1 << 24), b""):
            h.update(blk)
    assert h.hexdigest() == want


def test_c_example_builds_and_runs_through_the_abi(tmp_path):
    """examples/abi_example.c: the ABI used from plain C (gcc, include/rsgpu.h, -lrsgpu), no Python between."""
    from gpu_lib import ROOT
    exe = str(tmp_path / "abi_example")
    libdir = os.path.join(ROOT, "gpu-rscode_b200")
    subprocess.run(["gcc", "-O2", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "examples", "abi_example.c"),
                    "-L", libdir, "-lrsgpu", "-Wl,-rpath," + libdir, "-o", exe], check=True)
    for nbytes in ("1048576", "100003", "12582912"):       # aligned, ragged, long enough for the byte-table kernel
        r = subprocess.run([exe, nbytes], capture_output=True, text=True, timeout=120)
        assert r.returncode == 0, r.stdout + r.stderr
        assert "4 lost natives rebuilt, identical to the originals" in r.stdout
