"""ctypes access to the parity checker: oracle/liboracle.so (our C restatement) and, when it was
built in the container that has /root/reference, oracle/_ref/libcpurs_ref.so (the reference's own
src/cpu-rs.c, unmodified).  Test infrastructure only -- the product never imports this."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "liboracle.so")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libcpurs_ref.so")
REF_BIN = os.path.join(ORACLE_DIR, "_ref", "CPU-RS")

_u8p = C.POINTER(C.c_uint8)


def _p(a: np.ndarray):
    assert a.dtype == np.uint8 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_u8p)


def build_oracle():
    """(Re)build liboracle.so and, if the reference tree is here, oracle/_ref/."""
    subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True)


class Oracle:
    """Our restatement (prefix orc_) -- same call shapes as the reference functions."""

    def __init__(self):
        if not os.path.exists(ORACLE_SO):
            build_oracle()
        L = C.CDLL(ORACLE_SO)
        L.orc_gf_mul.restype = C.c_uint8
        L.orc_gf_mul.argtypes = [C.c_uint8, C.c_uint8]
        L.orc_gf_div.restype = C.c_uint8
        L.orc_gf_div.argtypes = [C.c_uint8, C.c_uint8]
        L.orc_gf_pow.restype = C.c_uint8
        L.orc_gf_pow.argtypes = [C.c_uint8, C.c_uint8]
        L.orc_log_table.restype = _u8p
        L.orc_alog_table.restype = _u8p
        L.orc_gen_encoding_matrix.argtypes = [_u8p, C.c_int, C.c_int]
        L.orc_gen_total_encoding_matrix.argtypes = [_u8p, C.c_int, C.c_int]
        L.orc_matrix_mul.argtypes = [_u8p, _u8p, _u8p, C.c_int, C.c_int, C.c_int]
        L.orc_invert_matrix.restype = C.c_int
        L.orc_invert_matrix.argtypes = [_u8p, _u8p, C.c_int]
        L.orc_encode_file.restype = C.c_int
        L.orc_encode_file.argtypes = [C.c_char_p, C.c_int, C.c_int]
        L.orc_decode_file.restype = C.c_int
        L.orc_decode_file.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.POINTER(C.c_int)]
        L.orc_setup_tables()
        self.L = L

    def log_table(self):
        return np.ctypeslib.as_array(self.L.orc_log_table(), shape=(256,)).copy()

    def alog_table(self):
        return np.ctypeslib.as_array(self.L.orc_alog_table(), shape=(256,)).copy()

    def gf_mul(self, a, b):
        return self.L.orc_gf_mul(a, b)

    def gf_div(self, a, b):
        return self.L.orc_gf_div(a, b)

    def gf_pow(self, a, e):
        return self.L.orc_gf_pow(a, e)

    def gen_encoding_matrix(self, p, k):
        V = np.zeros((p, k), np.uint8)
        self.L.orc_gen_encoding_matrix(_p(V), p, k)
        return V

    def gen_total_encoding_matrix(self, k, p):
        T = np.zeros((k + p, k), np.uint8)
        self.L.orc_gen_total_encoding_matrix(_p(T), k, p)
        return T

    def matrix_mul(self, A, B):
        """returns A (n x p) * B (p x m) over GF(2^8), starting from a zeroed C as the callers do."""
        A = np.ascontiguousarray(A, np.uint8)
        B = np.ascontiguousarray(B, np.uint8)
        n, p = A.shape
        p2, m = B.shape
        assert p == p2
        Cm = np.zeros((n, m), np.uint8)
        self.L.orc_matrix_mul(_p(A), _p(B), _p(Cm), n, p, m)
        return Cm

    def invert_matrix(self, M):
        """returns (inverse, status); status 0 = reference-correct, 1 = reference swapped columns
        (result is what the reference computes, not the inverse), 2 = reference would hit UB."""
        M = np.array(M, np.uint8, copy=True, order="C")
        n = M.shape[0]
        R = np.zeros((n, n), np.uint8)
        st = self.L.orc_invert_matrix(_p(M), _p(R), n)
        return R, st

    def encode_file(self, name, k, p):
        return self.L.orc_encode_file(name.encode(), k, p)

    def decode_file(self, in_file, conf, out):
        st = C.c_int(-1)
        rc = self.L.orc_decode_file(in_file.encode(), conf.encode(), out.encode() if out else None, C.byref(st))
        return rc, st.value


class Reference:
    """The reference's own cpu-rs.c as a shared object (globals live inside that .so)."""

    def __init__(self):
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(REF_SO)
        # RTLD_LOCAL (ctypes default): its gf_mul/matrix_mul/encode_file names stay private.
        L = C.CDLL(REF_SO)
        L.setup_tables.argtypes = [C.c_int]
        L.gf_mul.restype = C.c_uint8
        L.gf_mul.argtypes = [C.c_uint8, C.c_uint8]
        L.gf_div.restype = C.c_uint8
        L.gf_div.argtypes = [C.c_uint8, C.c_uint8]
        L.gf_pow.restype = C.c_uint8
        L.gf_pow.argtypes = [C.c_uint8, C.c_uint8]
        L.gen_encoding_matrix.argtypes = [_u8p, C.c_int, C.c_int]
        L.gen_total_encoding_matrix.argtypes = [_u8p, C.c_int, C.c_int]
        L.matrix_mul.argtypes = [_u8p, _u8p, _u8p, C.c_int, C.c_int, C.c_int]
        L.invert_matrix.argtypes = [_u8p, _u8p, C.c_int]
        L.setup_tables(8)
        self.L = L

    def log_table(self):
        return np.ctypeslib.as_array((C.c_uint8 * 256).in_dll(self.L, "gflog")).copy()

    def alog_table(self):
        return np.ctypeslib.as_array((C.c_uint8 * 256).in_dll(self.L, "gfilog")).copy()

    def gf_mul(self, a, b):
        return self.L.gf_mul(a, b)

    def gf_div(self, a, b):
        return self.L.gf_div(a, b)

    def gf_pow(self, a, e):
        return self.L.gf_pow(a, e)

    def gen_encoding_matrix(self, p, k):
        V = np.zeros((p, k), np.uint8)
        self.L.gen_encoding_matrix(_p(V), p, k)
        return V

    def gen_total_encoding_matrix(self, k, p):
        T = np.zeros((k + p, k), np.uint8)
        self.L.gen_total_encoding_matrix(_p(T), k, p)
        return T

    def matrix_mul(self, A, B):
        A = np.ascontiguousarray(A, np.uint8)
        B = np.ascontiguousarray(B, np.uint8)
        n, p = A.shape
        _, m = B.shape
        Cm = np.zeros((n, m), np.uint8)
        self.L.matrix_mul(_p(A), _p(B), _p(Cm), n, p, m)
        return Cm

    def invert_matrix(self, M):
        M = np.array(M, np.uint8, copy=True, order="C")
        n = M.shape[0]
        R = np.zeros((n, n), np.uint8)
        self.L.invert_matrix(_p(M), _p(R), n)
        return R


def have_reference():
    return os.path.exists(REF_SO) and os.path.exists(REF_BIN)


REF_GPU_SO = os.path.join(ORACLE_DIR, "_ref", "libgpurs_ref.so")


def have_reference_gpu():
    return os.path.exists(REF_GPU_SO)


class ReferenceGPU:
    """The reference's own CUDA launchers (src/matrix.cu:752,767,838), recompiled for sm_100a behind
    oracle/ref_gpu_shim.cu.  Device pointers are plain integers (torch tensors' data_ptr())."""

    MAX_GRID = 2147483647   # what src/encode.cu:351-355 falls back to without -p

    def __init__(self):
        if not os.path.exists(REF_GPU_SO):
            raise FileNotFoundError(REF_GPU_SO)
        L = C.CDLL(REF_GPU_SO)
        for name in ("refgpu_encode_chunk", "refgpu_decode_chunk"):
            f = getattr(L, name)
            f.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
            f.restype = C.c_float
        L.refgpu_gen_encoding_matrix.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.refgpu_gen_encoding_matrix.restype = C.c_int
        self.L = L

    def encode_chunk(self, data_d, coeff_d, code_d, k, p, chunk, grid_x=MAX_GRID):
        assert k * chunk < 2 ** 31, "the reference indexes with int"
        return float(self.L.refgpu_encode_chunk(data_d, coeff_d, code_d, k, p, chunk, grid_x))

    def decode_chunk(self, data_d, inv_d, code_d, k, p, chunk, grid_x=MAX_GRID):
        assert k * chunk < 2 ** 31, "the reference indexes with int"
        return float(self.L.refgpu_decode_chunk(data_d, inv_d, code_d, k, p, chunk, grid_x))

    def gen_encoding_matrix(self, matrix_d, p, k):
        rc = self.L.refgpu_gen_encoding_matrix(matrix_d, p, k)
        if rc:
            raise RuntimeError(f"reference gen_encoding_matrix: cuda error {rc}")
