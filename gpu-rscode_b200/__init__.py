"""gpu-rscode_b200 -- Python face of the B200-native Reed-Solomon engine.

This package is glue only: it loads the in-tree ``librsgpu.so`` (hand-written sm_100a CUDA kernels
behind the C ABI of ``include/rsgpu.h``) with ctypes and mirrors that ABI one to one.  There is no
Python/torch/CPU implementation of the coding path here: if the shared library is missing or the
GPU is unusable every call raises.  (The CPU oracle under ``oracle/`` is test infrastructure and is
never imported from here.)

The directory name carries a hyphen, so import it with
``importlib.import_module("gpu-rscode_b200")`` (the repo root must be on ``sys.path``).
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RSGPU_LIB") or os.path.join(HERE, "librsgpu.so")   # RSGPU_LIB: A/B builds (developer)
HOST_LIB_PATH = os.path.join(HERE, "librs_host.so")
RS_BIN = os.path.join(HERE, "RS")

RSGPU_OK = 0
RSGPU_EINVAL = -1
RSGPU_ECUDA = -2
RSGPU_ESINGULAR = -3
RSGPU_ENOMEM = -4
RSGPU_ESTATE = -5
RSGPU_ETOOBIG = -6

# every symbol include/rsgpu.h declares (tests check the library exports each of them)
ABI_SYMBOLS = [
    "rsgpu_strerror", "rsgpu_last_error", "rsgpu_device_count", "rsgpu_device_max_grid", "rsgpu_create", "rsgpu_create_with_matrix",
    "rsgpu_destroy", "rsgpu_k", "rsgpu_p", "rsgpu_device", "rsgpu_get_total_matrix", "rsgpu_encode",
    "rsgpu_set_survivors", "rsgpu_get_decode_matrix", "rsgpu_get_erased", "rsgpu_decode", "rsgpu_decode_erased",
    "rsgpu_invert_matrix", "rsgpu_matmul", "rsgpu_encode_host", "rsgpu_decode_host", "rsgpu_decode_erased_host", "rsgpu_host_alloc",
    "rsgpu_host_free", "rsgpu_host_register", "rsgpu_host_unregister", "rsgpu_dev_alloc", "rsgpu_dev_free",
    "rsgpu_memcpy_h2d", "rsgpu_memcpy_d2h", "rsgpu_sync", "rsgpu_set_tuning", "rsgpu_launch_count",
    "rsgpu_encode_noalloc", "rsgpu_create_generator", "rsgpu_set_repair", "rsgpu_repair", "rsgpu_enable_peer", "rsgpu_encode_rows",
    "rsgpu_encode_batch", "rsgpu_decode_erased_batch",
]


class RSGPUError(RuntimeError):
    def __init__(self, code, detail=""):
        self.code = code
        super().__init__(f"rsgpu error {code}: {detail}")


class SingularError(RSGPUError):
    pass


def build(verbose=False):
    """Compile librsgpu.so (+ host layer) in-tree for sm_100a with the Makefile next to this file."""
    subprocess.run(["make", "-C", HERE, "-j8"] + ([] if verbose else ["-s"]), check=True)


_lib = None


def lib():
    """The loaded C ABI.  Raises if the CUDA library has not been built -- there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RSGPUError(RSGPU_ESTATE, f"{LIB_PATH} is missing: run `make -C gpu-rscode_b200` "
                                       "(or __graft_entry__.build()); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, sz, i32, u8p = C.c_void_p, C.c_size_t, C.c_int, C.c_void_p
    L.rsgpu_strerror.restype = C.c_char_p
    L.rsgpu_strerror.argtypes = [i32]
    L.rsgpu_last_error.restype = C.c_char_p
    L.rsgpu_device_count.argtypes = [C.POINTER(i32)]
    L.rsgpu_device_max_grid.argtypes = [i32, C.POINTER(i32)]
    L.rsgpu_create.argtypes = [C.POINTER(vp), i32, i32, i32]
    L.rsgpu_create_with_matrix.argtypes = [C.POINTER(vp), i32, i32, i32, u8p]
    L.rsgpu_create_generator.argtypes = [C.POINTER(vp), i32, i32, i32, i32]
    L.rsgpu_set_repair.argtypes = [vp, C.POINTER(i32), C.POINTER(i32), i32]
    L.rsgpu_repair.argtypes = [vp, u8p, sz, u8p, sz, sz, vp]
    L.rsgpu_enable_peer.argtypes = [vp, i32]
    L.rsgpu_encode_rows.argtypes = [vp, u8p, sz, C.POINTER(vp), sz, vp]
    L.rsgpu_encode_batch.argtypes = [vp, u8p, sz, sz, u8p, sz, sz, sz, sz, vp]
    L.rsgpu_decode_erased_batch.argtypes = [vp, u8p, sz, sz, u8p, sz, sz, sz, sz, vp]
    L.rsgpu_destroy.argtypes = [vp]
    L.rsgpu_destroy.restype = None
    for f in ("rsgpu_k", "rsgpu_p", "rsgpu_device"):
        getattr(L, f).argtypes = [vp]
    L.rsgpu_get_total_matrix.argtypes = [vp, u8p]
    L.rsgpu_encode.argtypes = [vp, u8p, sz, u8p, sz, sz, vp]
    L.rsgpu_set_survivors.argtypes = [vp, C.POINTER(i32)]
    L.rsgpu_get_decode_matrix.argtypes = [vp, u8p]
    L.rsgpu_get_erased.argtypes = [vp, C.POINTER(i32), C.POINTER(i32)]
    L.rsgpu_decode.argtypes = [vp, u8p, sz, u8p, sz, sz, vp]
    L.rsgpu_decode_erased.argtypes = [vp, u8p, sz, u8p, sz, sz, vp]
    L.rsgpu_invert_matrix.argtypes = [vp, u8p, u8p, i32]
    L.rsgpu_matmul.argtypes = [vp, u8p, i32, u8p, sz, u8p, sz, sz, vp]
    L.rsgpu_encode_host.argtypes = [vp, u8p, sz, u8p, sz, sz, i32]
    L.rsgpu_decode_host.argtypes = [vp, u8p, sz, u8p, sz, sz, i32]
    L.rsgpu_decode_erased_host.argtypes = [vp, u8p, sz, u8p, sz, sz, i32]
    L.rsgpu_host_alloc.argtypes = [C.POINTER(vp), sz]
    L.rsgpu_host_free.argtypes = [vp]
    L.rsgpu_host_register.argtypes = [vp, sz]
    L.rsgpu_host_unregister.argtypes = [vp]
    L.rsgpu_dev_alloc.argtypes = [vp, C.POINTER(vp), sz]
    L.rsgpu_dev_free.argtypes = [vp, vp]
    L.rsgpu_memcpy_h2d.argtypes = [vp, vp, vp, sz]
    L.rsgpu_memcpy_d2h.argtypes = [vp, vp, vp, sz]
    L.rsgpu_sync.argtypes = [vp]
    L.rsgpu_set_tuning.argtypes = [vp, i32, i32, i32]
    L.rsgpu_launch_count.argtypes = [vp]
    L.rsgpu_launch_count.restype = C.c_longlong
    L.rsgpu_encode_noalloc.argtypes = [i32, i32, u8p, sz, u8p, sz, sz, vp]
    _lib = L
    return L


def _check(rc):
    if rc == RSGPU_OK:
        return
    L = lib()
    detail = L.rsgpu_strerror(rc).decode() + ": " + L.rsgpu_last_error().decode()
    if rc == RSGPU_ESINGULAR:
        raise SingularError(rc, detail)
    raise RSGPUError(rc, detail)


def device_count():
    n = C.c_int(0)
    _check(lib().rsgpu_device_count(C.byref(n)))
    return n.value


def _ptr(x):
    """Raw address of a numpy array (host) / torch tensor (device or host) / int."""
    if isinstance(x, int):
        return x
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    if hasattr(x, "data_ptr"):
        return x.data_ptr()
    raise TypeError(type(x))


class RSContext:
    """One RS(k, k+p) coding context on one GPU -- a thin mirror of the rsgpu_* C ABI.

    Pointer arguments are raw addresses (ints), numpy arrays (host entry points) or torch tensors
    (device entry points); pitches and lengths are bytes.  `stream` is a raw cudaStream_t (int), e.g.
    ``torch.cuda.current_stream().cuda_stream``.
    """

    def __init__(self, k, p, device=0, total_matrix=None, generator=None):
        self._h = C.c_void_p(None)
        L = lib()
        if generator is not None:
            kind = {"vandermonde": 0, "cauchy": 1}[generator] if isinstance(generator, str) else int(generator)
            _check(L.rsgpu_create_generator(C.byref(self._h), device, k, p, kind))
        elif total_matrix is None:
            _check(L.rsgpu_create(C.byref(self._h), device, k, p))
        else:
            tm = np.ascontiguousarray(total_matrix, np.uint8)
            assert tm.shape == (k + p, k)
            _check(L.rsgpu_create_with_matrix(C.byref(self._h), device, k, p, tm.ctypes.data))
        self.k, self.p, self.device = k, p, device

    def close(self):
        if self._h:
            lib().rsgpu_destroy(self._h)
            self._h = C.c_void_p(None)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- matrices ---------------------------------------------------------------------------
    def total_matrix(self):
        T = np.zeros((self.k + self.p, self.k), np.uint8)
        _check(lib().rsgpu_get_total_matrix(self._h, T.ctypes.data))
        return T

    def generator(self):
        return self.total_matrix()[self.k:]

    def set_survivors(self, idx):
        arr = (C.c_int * self.k)(*[int(i) for i in idx])
        _check(lib().rsgpu_set_survivors(self._h, arr))

    def decode_matrix(self):
        M = np.zeros((self.k, self.k), np.uint8)
        _check(lib().rsgpu_get_decode_matrix(self._h, M.ctypes.data))
        return M

    def erased(self):
        arr = (C.c_int * 256)()
        n = C.c_int(0)
        _check(lib().rsgpu_get_erased(self._h, arr, C.byref(n)))
        return [arr[i] for i in range(n.value)]

    def set_repair(self, survivors, wanted):
        sv = (C.c_int * self.k)(*[int(i) for i in survivors])
        wt = (C.c_int * len(wanted))(*[int(i) for i in wanted])
        _check(lib().rsgpu_set_repair(self._h, sv, wt, len(wanted)))

    def repair(self, surv_d, surv_pitch, out_d, out_pitch, length, stream=0):
        _check(lib().rsgpu_repair(self._h, _ptr(surv_d), surv_pitch, _ptr(out_d), out_pitch, length, stream))

    def invert_matrix(self, M):
        M = np.ascontiguousarray(M, np.uint8)
        n = M.shape[0]
        assert M.shape == (n, n)
        inv = np.zeros((n, n), np.uint8)
        _check(lib().rsgpu_invert_matrix(self._h, M.ctypes.data, inv.ctypes.data, n))
        return inv

    # -- device-pointer coding ----------------------------------------------------------------
    def encode(self, data_d, data_pitch, parity_d, parity_pitch, length, stream=0):
        _check(lib().rsgpu_encode(self._h, _ptr(data_d), data_pitch, _ptr(parity_d), parity_pitch, length, stream))

    def encode_batch(self, data_d, data_pitch, data_stripe_pitch, parity_d, parity_pitch, parity_stripe_pitch, length,
                     n_stripes, stream=0):
        """n_stripes stripes in one call: fragment j of stripe s at data_d + s*data_stripe_pitch + j*data_pitch."""
        _check(lib().rsgpu_encode_batch(self._h, _ptr(data_d), data_pitch, data_stripe_pitch, _ptr(parity_d), parity_pitch,
                                        parity_stripe_pitch, length, n_stripes, stream))

    def decode_erased_batch(self, surv_d, surv_pitch, surv_stripe_pitch, out_d, out_pitch, out_stripe_pitch, length,
                            n_stripes, stream=0):
        _check(lib().rsgpu_decode_erased_batch(self._h, _ptr(surv_d), surv_pitch, surv_stripe_pitch, _ptr(out_d), out_pitch,
                                               out_stripe_pitch, length, n_stripes, stream))

    def enable_peer(self, peer_device):
        _check(lib().rsgpu_enable_peer(self._h, peer_device))

    def encode_rows(self, data_d, data_pitch, parity_rows, length, stream=0):
        """parity_rows: p device pointers / tensors, each may live on a peer GPU (after enable_peer)."""
        arr = (C.c_void_p * self.p)(*[_ptr(r) for r in parity_rows])
        _check(lib().rsgpu_encode_rows(self._h, _ptr(data_d), data_pitch, arr, length, stream))

    def decode(self, surv_d, surv_pitch, natives_d, natives_pitch, length, stream=0):
        _check(lib().rsgpu_decode(self._h, _ptr(surv_d), surv_pitch, _ptr(natives_d), natives_pitch, length, stream))

    def decode_erased(self, surv_d, surv_pitch, out_d, out_pitch, length, stream=0):
        _check(lib().rsgpu_decode_erased(self._h, _ptr(surv_d), surv_pitch, _ptr(out_d), out_pitch, length, stream))

    def matmul(self, matrix, in_d, in_pitch, out_d, out_pitch, length, stream=0):
        M = np.ascontiguousarray(matrix, np.uint8)
        assert M.ndim == 2 and M.shape[1] == self.k
        _check(lib().rsgpu_matmul(self._h, M.ctypes.data, M.shape[0], _ptr(in_d), in_pitch, _ptr(out_d), out_pitch,
                                  length, stream))

    # -- host-pointer coding (pipelined H2D / kernel / D2H) --------------------------------------
    def encode_host(self, data, data_pitch, parity, parity_pitch, length, slots=0):
        _check(lib().rsgpu_encode_host(self._h, _ptr(data), data_pitch, _ptr(parity), parity_pitch, length, slots))

    def decode_host(self, surv, surv_pitch, natives, natives_pitch, length, slots=0):
        _check(lib().rsgpu_decode_host(self._h, _ptr(surv), surv_pitch, _ptr(natives), natives_pitch, length, slots))

    def decode_erased_host(self, surv, surv_pitch, out, out_pitch, length, slots=0):
        _check(lib().rsgpu_decode_erased_host(self._h, _ptr(surv), surv_pitch, _ptr(out), out_pitch, length, slots))

    # -- misc -------------------------------------------------------------------------------------
    def set_tuning(self, variant=0, threads=0, max_blocks=0):
        _check(lib().rsgpu_set_tuning(self._h, variant, threads, max_blocks))

    def launch_count(self):
        return int(lib().rsgpu_launch_count(self._h))

    def sync(self):
        _check(lib().rsgpu_sync(self._h))

    def dev_alloc(self, nbytes):
        p = C.c_void_p(None)
        _check(lib().rsgpu_dev_alloc(self._h, C.byref(p), nbytes))
        return p.value

    def dev_free(self, ptr):
        _check(lib().rsgpu_dev_free(self._h, ptr))

    def h2d(self, dst_d, src: np.ndarray):
        _check(lib().rsgpu_memcpy_h2d(self._h, dst_d, src.ctypes.data, src.nbytes))

    def d2h(self, dst: np.ndarray, src_d):
        _check(lib().rsgpu_memcpy_d2h(self._h, dst.ctypes.data, src_d, dst.nbytes))


def encode_noalloc(k, p, data_d, data_pitch, parity_d, parity_pitch, length, stream=0):
    """rsgpu_encode_noalloc on the current CUDA device: no context, no allocation (k <= 32, p <= 8)."""
    _check(lib().rsgpu_encode_noalloc(k, p, _ptr(data_d), data_pitch, _ptr(parity_d), parity_pitch, length, stream))


def host_alloc(nbytes):
    """Pinned host buffer as a numpy uint8 array (freed with host_free(arr))."""
    p = C.c_void_p(None)
    _check(lib().rsgpu_host_alloc(C.byref(p), nbytes))
    arr = np.ctypeslib.as_array((C.c_uint8 * nbytes).from_address(p.value))
    arr._rsgpu_ptr = p.value if hasattr(arr, "__dict__") else None
    return arr, p.value


def host_free(ptr):
    _check(lib().rsgpu_host_free(ptr))
