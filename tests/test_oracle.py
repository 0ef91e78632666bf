"""CPU tests that PIN the oracle (oracle/rs_oracle.c) before anything else trusts it.

Three anchors: (1) known-answer vectors in the reference tree, (2) tests/golden/golden.json, produced
by running the reference's own cpu-rs.c (tests/golden/make_golden.py), (3) the reference compiled
unmodified into oracle/_ref/ when present (always in the build container; it also travels to the GPU
box, but nothing here needs /root/reference at run time).
"""
import hashlib
import json
import os
import shutil
import subprocess

import numpy as np
import pytest

from oracle_lib import REF_BIN, Oracle, Reference, have_reference
from rsdata import pattern_bytes, splitmix64_bytes

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "golden.json")))


def sha(b):
    return hashlib.sha256(bytes(b)).hexdigest()


@pytest.fixture(scope="module")
def orc():
    return Oracle()


@pytest.fixture(scope="module")
def ref():
    if not have_reference():
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    return Reference()


def make_input(spec):
    if spec["gen"] == "pattern":
        return pattern_bytes(spec["mul"], spec["add"], spec["nbytes"])
    return splitmix64_bytes(spec["seed"], spec["nbytes"])


# ---- (1) known answers held by the reference tree ----------------------------------------------

def test_splitmix_first_bytes():
    assert splitmix64_bytes(1, 16).tobytes().hex() == "c15c0289ec2d0a9167ec8e65a18debbe"


def test_exp_table_literal(orc):
    # src/matrix.cu:36 gfexp_cMem starts 1,2,4,8,16,32,64,128,29,58,116,232,205,135,19,38,76,152,45,90,...
    head = [1, 2, 4, 8, 16, 32, 64, 128, 29, 58, 116, 232, 205, 135, 19, 38, 76, 152, 45, 90, 180, 117, 234, 201, 143, 3, 6]
    assert list(orc.alog_table()[: len(head)]) == head


def test_log_table_literal(orc):
    # src/matrix.cu:37 gflog_cMem: 510(for 0),0,1,25,2,50,26,198,3,223,51,238,27,104,199,75,4,100,224,14,52,141,239,129,28,193
    head = [0, 1, 25, 2, 50, 26, 198, 3, 223, 51, 238, 27, 104, 199, 75, 4, 100, 224, 14, 52, 141, 239, 129, 28, 193]
    assert list(orc.log_table()[1 : 1 + len(head)]) == head


def test_inverse_kat_slide(orc):
    # doc/slide.tex:501-557: [1000;0100;1234;1111] -> rows 104 187 186 210 / 105 186 186 211
    M = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [1, 2, 3, 4], [1, 1, 1, 1]], np.uint8)
    inv, st = orc.invert_matrix(M)
    assert st == 0
    assert inv.tolist() == [[1, 0, 0, 0], [0, 1, 0, 0], [104, 187, 186, 210], [105, 186, 186, 211]]


def test_inverse_kat_decode_gj(orc):
    # src/decode-gj.cu:1282 testMatrix and :1453 its expected inverse
    M = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [1, 1, 1, 1], [1, 2, 3, 4]], np.uint8)
    inv, st = orc.invert_matrix(M)
    assert st == 0
    assert inv.tolist() == [[1, 0, 0, 0], [0, 1, 0, 0], [104, 187, 210, 186], [105, 186, 211, 186]]


# ---- (2) golden fixtures generated from the reference -------------------------------------------

def test_tables_vs_golden(orc):
    assert orc.log_table().tobytes().hex() == GOLD["log"]
    assert orc.alog_table().tobytes().hex() == GOLD["alog"]


def test_gf_ops_exhaustive_vs_golden(orc):
    mul = np.array([[orc.gf_mul(a, b) for b in range(256)] for a in range(256)], np.uint8)
    div = np.array([[orc.gf_div(a, b) for b in range(256)] for a in range(256)], np.uint8)
    pw = np.array([[orc.gf_pow(a, b) for b in range(256)] for a in range(256)], np.uint8)
    assert sha(mul.tobytes()) == GOLD["gf_mul_sha256"]
    assert sha(div.tobytes()) == GOLD["gf_div_sha256"]
    assert sha(pw.tobytes()) == GOLD["gf_pow_sha256"]
    for a, row in GOLD["gf_mul_rows"].items():
        assert mul[int(a)].tobytes().hex() == row


def test_gf_field_axioms(orc):
    # independent of the reference: a*b computed by shift-and-reduce with 0x11D
    def slow(a, b):
        r = 0
        while b:
            if b & 1:
                r ^= a
            a = ((a << 1) ^ 0x11D) if a & 0x80 else (a << 1)
            b >>= 1
        return r
    for a in range(0, 256, 7):
        for b in range(256):
            assert orc.gf_mul(a, b) == slow(a, b)
    for a in range(1, 256):
        assert orc.gf_mul(a, orc.gf_div(1, a)) == 1


@pytest.mark.parametrize("key", sorted(GOLD["vandermonde"].keys()))
def test_vandermonde_vs_golden(orc, key):
    k, p = map(int, key.split(","))
    assert orc.gen_encoding_matrix(p, k).tobytes().hex() == GOLD["vandermonde"][key]


def test_vandermonde_survey_rows(orc):
    V = orc.gen_encoding_matrix(4, 10)
    assert V[2].tolist() == [1, 4, 5, 16, 17, 20, 21, 64, 65, 68]
    assert V[3].tolist() == [1, 8, 15, 64, 85, 120, 107, 58, 115, 146]
    V = orc.gen_encoding_matrix(8, 32)
    assert V[7].tolist() == [1, 128, 255, 19, 226, 98, 206, 117, 192, 68, 79, 87, 43, 199, 38, 24, 174, 148, 67,
                             189, 138, 84, 33, 200, 30, 36, 83, 51, 93, 234, 6, 156]


@pytest.mark.parametrize("case", GOLD["matmul"], ids=lambda c: f"k{c['k']}p{c['p']}m{c['m']}")
def test_matmul_vs_golden(orc, case):
    k, p, m = case["k"], case["p"], case["m"]
    data = splitmix64_bytes(case["seed"], k * m).reshape(k, m)
    code = orc.matrix_mul(orc.gen_encoding_matrix(p, k), data)
    assert sha(code.tobytes()) == case["code_sha256"]


def test_matmul_accumulates(orc):
    # cpu-rs.c:194 uses ^=, so a pre-filled C is XORed into, not overwritten
    import ctypes as C
    A = np.array([[3, 7]], np.uint8)
    B = np.array([[5, 9, 0], [1, 2, 255]], np.uint8)
    base = orc.matrix_mul(A, B)
    Cm = np.array([[0xAA, 0x55, 0x0F]], np.uint8)
    u8p = C.POINTER(C.c_uint8)
    orc.L.orc_matrix_mul(A.ctypes.data_as(u8p), B.ctypes.data_as(u8p), Cm.ctypes.data_as(u8p), 1, 2, 3)
    assert (Cm == (base ^ np.array([[0xAA, 0x55, 0x0F]], np.uint8))).all()


@pytest.mark.parametrize("case", GOLD["inverse"] + GOLD["inverse_natural_rs46"],
                         ids=lambda c: f"k{c['k']}-{'_'.join(map(str, c['order'][:5]))}")
def test_inverse_vs_golden(orc, case):
    k, p = case["k"], case["p"]
    T = orc.gen_total_encoding_matrix(k, p)
    sub = T[case["order"]]
    inv, st = orc.invert_matrix(sub)
    assert inv.tobytes().hex() == case["inverse"]
    ident = orc.matrix_mul(sub, inv)
    is_inv = bool((ident == np.eye(k, dtype=np.uint8)).all())
    if "is_true_inverse" in case:
        assert is_inv == case["is_true_inverse"]
    if st == 0:
        assert is_inv


@pytest.mark.parametrize("ci", range(len(GOLD["encode"])))
def test_encode_file_vs_golden(orc, ci, tmp_path, monkeypatch):
    spec = GOLD["encode"][ci]
    if spec["nbytes"] > 2_000_000:
        pytest.skip("large seeded case is covered by test_encode_large_vs_golden")
    data = make_input(spec)
    assert sha(data.tobytes()) == spec["input_sha256"]
    monkeypatch.chdir(tmp_path)
    (tmp_path / "f.bin").write_bytes(data.tobytes())
    assert orc.encode_file("f.bin", spec["k"], spec["p"]) == 0
    assert (tmp_path / "f.bin.METADATA").read_text() == spec["metadata"]
    for i, want in enumerate(spec["fragment_sha256"]):
        frag = (tmp_path / f"_{i}_f.bin").read_bytes()
        assert len(frag) == spec["chunk"]
        assert sha(frag) == want, f"fragment {i}"
        if "fragment_hex" in spec:
            assert frag.hex() == spec["fragment_hex"][i]


@pytest.mark.parametrize("ci", [i for i, s in enumerate(GOLD["encode"]) if s["nbytes"] > 2_000_000])
def test_encode_large_vs_golden(orc, ci):
    # SURVEY 8c seeded fixtures (10,485,773 B RS(10,14) and 4 MiB RS(4,6)): parity via matrix_mul
    spec = GOLD["encode"][ci]
    k, p, L = spec["k"], spec["p"], spec["chunk"]
    data = np.zeros(k * L, np.uint8)
    data[: spec["nbytes"]] = make_input(spec)
    code = orc.matrix_mul(orc.gen_encoding_matrix(p, k), data.reshape(k, L))
    for i in range(p):
        assert sha(code[i].tobytes()) == spec["fragment_sha256"][k + i]


@pytest.mark.parametrize("di", range(len(GOLD["decode"])))
def test_decode_file_vs_golden(orc, di, tmp_path, monkeypatch):
    case = GOLD["decode"][di]
    spec = GOLD["encode"][case["encode_case"]]
    data = make_input(spec)
    monkeypatch.chdir(tmp_path)
    (tmp_path / "f.bin").write_bytes(data.tobytes())
    assert orc.encode_file("f.bin", spec["k"], spec["p"]) == 0
    (tmp_path / "conf").write_text("".join(f"_{i}_f.bin\n" for i in case["order"]))
    rc, st = orc.decode_file("f.bin", "conf", "out.bin")
    assert rc == 0
    got = (tmp_path / "out.bin").read_bytes()
    # bit-identical to what the reference wrote, including the orders where the reference is wrong
    assert sha(got) == case["output_sha256"]
    assert (got == data.tobytes()) == case["reference_correct"]
    assert (st == 0) == case["reference_correct"]


# ---- (3) the reference itself, when built ---------------------------------------------------------

def test_oracle_vs_reference_lib(orc, ref):
    assert (orc.log_table() == ref.log_table()).all()
    assert (orc.alog_table() == ref.alog_table()).all()
    rng = np.random.default_rng(5)
    for k, p in [(4, 2), (10, 4), (16, 4), (32, 8), (3, 5), (1, 1), (255, 1)]:
        assert (orc.gen_encoding_matrix(p, k) == ref.gen_encoding_matrix(p, k)).all()
        m = int(rng.integers(1, 3000))
        data = rng.integers(0, 256, (k, m), dtype=np.uint8)
        A = rng.integers(0, 256, (p, k), dtype=np.uint8)
        assert (orc.matrix_mul(A, data) == ref.matrix_mul(A, data)).all()


def test_oracle_inverse_vs_reference_lib_random_orders(orc, ref):
    # every survivor ORDER, also those where the reference swaps columns (its wrong answers included)
    rng = np.random.default_rng(9)
    for k, p in [(4, 2), (10, 4), (16, 4)]:
        T = orc.gen_total_encoding_matrix(k, p)
        n_checked = 0
        for _ in range(300):
            order = rng.permutation(k + p)[:k]
            sub = T[order]
            inv_o, st = orc.invert_matrix(sub)
            if st == 2:
                continue  # reference would index column -1: undefined, not comparable
            inv_r = ref.invert_matrix(sub)
            assert (inv_o == inv_r).all(), (k, p, order.tolist(), st)
            n_checked += 1
        assert n_checked > 100


def test_reference_binary_roundtrip_matches_oracle(orc, ref, tmp_path):
    data = splitmix64_bytes(77, 12345)
    for who in ("ref", "orc"):
        d = tmp_path / who
        d.mkdir()
        (d / "g.bin").write_bytes(data.tobytes())
    subprocess.run([REF_BIN, "-k", "5", "-n", "8", "-e", "g.bin"], cwd=tmp_path / "ref", check=True,
                   stdout=subprocess.DEVNULL)
    cwd = os.getcwd()
    os.chdir(tmp_path / "orc")
    try:
        assert orc.encode_file("g.bin", 5, 3) == 0
    finally:
        os.chdir(cwd)
    for i in range(8):
        assert (tmp_path / "ref" / f"_{i}_g.bin").read_bytes() == (tmp_path / "orc" / f"_{i}_g.bin").read_bytes()
    assert (tmp_path / "ref" / "g.bin.METADATA").read_text() == (tmp_path / "orc" / "g.bin.METADATA").read_text()
