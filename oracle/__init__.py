"""ctypes front-end to the parity oracle.  TEST INFRASTRUCTURE ONLY.

Importable from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  Nothing under hpc_b200/ may import this package.

  oracle.lib()  -> liboracle.so   (CPU restatement; oracle/oracle.c)
  oracle.ref()  -> oracle/_ref/libcuxref.so (the reference's own host code), or None
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None
_ref = None
_ref_tried = False

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_f64p = C.POINTER(C.c_double)
_u16p = np.ctypeslib.ndpointer(np.uint16, flags="C_CONTIGUOUS")
_i8p = np.ctypeslib.ndpointer(np.int8, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_ll = C.c_longlong
_i = C.c_int
_f = C.c_float


def lib():
    global _lib
    if _lib is not None:
        return _lib
    path = os.path.join(_HERE, "liboracle.so")
    if not os.path.exists(path):
        from . import build
        build.build_oracle()
    L = C.CDLL(path)
    L.orc_gemm_f32_v0.argtypes = [_i, _i, _i, _f, _f32p, _i, _f32p, _i, _f, _f32p, _i]
    L.orc_gemm_f32_ikj.argtypes = [_i, _i, _i, _f, _f32p, _i, _f32p, _i, _f, _f32p, _i, _i]
    L.orc_gemm_f32_plain.argtypes = [_i, _i, _i, _f32p, _i, _f32p, _i, _f32p, _i]
    L.orc_hgemm_f16.argtypes = [_i, _i, _i, _f, _u16p, _i, _u16p, _i, _f, C.c_void_p, _i, _i, _i]
    L.orc_linear_s8.argtypes = [_i, _i, _i, _i8p, _i8p, _i32p, _i]
    L.orc_linear_bf16.argtypes = [_i, _i, _i, _u16p, _u16p, C.c_void_p, _i, _i]
    L.orc_linear_f32.argtypes = [_i, _i, _i, _f32p, _f32p, _f32p]
    L.orc_sum_f32.argtypes = [_ll, _f32p, _f32p]
    L.orc_dot_f32.argtypes = [_ll, _f32p, _f32p, _f32p]
    L.orc_dot_f32_v1.argtypes = [_ll, _f32p, _f32p, _f32p]
    L.orc_nrm2_f32.argtypes = [_ll, _f32p, _f32p]
    L.orc_dot_f64.argtypes = [_ll, _f32p, C.c_void_p, _f64p, _f64p]
    L.orc_scan_f32.argtypes = [_ll, _f32p, _f32p]
    L.orc_scan_i32.argtypes = [_ll, _i32p, _i32p, C.c_int32]
    L.orc_sum_i32.argtypes = [_ll, _i32p, _i32p]
    L.orc_cux_check_f32.argtypes = [_f32p, _f32p, _ll, _f]
    L.orc_cux_check_f32.restype = _f
    for n in ("orc_cvt_f32_to_f16", "orc_cvt_f32_to_bf16"):
        getattr(L, n).argtypes = [_ll, _f32p, _u16p]
    for n in ("orc_cvt_f16_to_f32", "orc_cvt_bf16_to_f32"):
        getattr(L, n).argtypes = [_ll, _u16p, _f32p]
    L.orc_num_threads.restype = _i
    _lib = L
    return L


def _bind_ref(path):
    R = C.CDLL(path)
    R.ref_cux_gemm.argtypes = [_i, _i, _i, _i, _f, _f32p, _i, _f32p, _i, _f, _f32p, _i]
    R.ref_cux_dot_v0.argtypes = [_i, _f32p, _f32p, _f32p]
    R.ref_cux_nrm2_v0.argtypes = [_i, _f32p, _f32p]
    if hasattr(R, "ref_std_dot_mt"):
        R.ref_std_dot_mt.argtypes = [_i, _f32p, _f32p, _i, _f32p]
    R.ref_reduce_host.argtypes = [_f32p, _f32p, _i]
    R.ref_hgemm_host.argtypes = [_i, _i, _i, _u16p, _i, _u16p, _i, _u16p, _i]
    R.ref_omp_gemm.argtypes = [_i, _i, _i, _i, _f, _f32p, _i, _f32p, _i, _f32p, _i]
    R.ref_linear_s8.argtypes = [_i, _i, _i, _i, _i8p, _i8p, _i32p]
    R.ref_linear_bf16.argtypes = [_i, _i, _i, _i, _u16p, _u16p, _u16p]
    R.ref_linear_f32.argtypes = [_i, _i, _i, _f32p, _f32p, _f32p]
    R.ref_scan_f32.argtypes = [_i, _f32p, _i, _f32p]
    return R


def ref():
    """The reference's own host functions, PARITY build (None when oracle/_ref was never built)."""
    global _ref, _ref_tried
    if _ref_tried:
        return _ref
    _ref_tried = True
    path = os.path.join(_HERE, "_ref", "libcuxref.so")
    if not os.path.exists(path):
        return None
    _ref = _bind_ref(path)
    return _ref


_ref_fast = None


def ref_fast():
    """The same functions from the -O3 build (oracle/build.py): for TIMING only -- floating-point
    contraction is allowed there, so nothing is ever compared against it.  Falls back to ref()."""
    global _ref_fast
    if _ref_fast is None:
        path = os.path.join(_HERE, "_ref", "libcuxref_fast.so")
        _ref_fast = _bind_ref(path) if os.path.exists(path) else ref()
    return _ref_fast


# --------------------------------------------------------------------------
# numpy-level helpers (what the tests call)
# --------------------------------------------------------------------------
def f32_to_f16(x):
    x = np.ascontiguousarray(x, np.float32)
    out = np.empty(x.shape, np.uint16)
    lib().orc_cvt_f32_to_f16(x.size, x.reshape(-1), out.reshape(-1))
    return out


def f16_to_f32(h):
    h = np.ascontiguousarray(h, np.uint16)
    out = np.empty(h.shape, np.float32)
    lib().orc_cvt_f16_to_f32(h.size, h.reshape(-1), out.reshape(-1))
    return out


def f32_to_bf16(x):
    x = np.ascontiguousarray(x, np.float32)
    out = np.empty(x.shape, np.uint16)
    lib().orc_cvt_f32_to_bf16(x.size, x.reshape(-1), out.reshape(-1))
    return out


def bf16_to_f32(h):
    h = np.ascontiguousarray(h, np.uint16)
    out = np.empty(h.shape, np.float32)
    lib().orc_cvt_bf16_to_f32(h.size, h.reshape(-1), out.reshape(-1))
    return out


def gemm_f32(A, B, C, alpha=1.0, beta=0.0, order="v0", threads=1):
    """cux gemm semantics; returns a new C.  order 'v0' = GemmHostV0, 'ikj' = GemmHostV1."""
    M, K = A.shape
    K2, N = B.shape
    assert K == K2 and C.shape == (M, N)
    out = np.array(C, np.float32, copy=True, order="C")
    A = np.ascontiguousarray(A, np.float32)
    B = np.ascontiguousarray(B, np.float32)
    if order == "v0":
        lib().orc_gemm_f32_v0(M, N, K, alpha, A, K, B, N, beta, out, N)
    else:
        lib().orc_gemm_f32_ikj(M, N, K, alpha, A, K, B, N, beta, out, N, threads)
    return out


def hgemm_f16(A16, B16, C=None, alpha=1.0, beta=0.0, c_f32=False, threads=1):
    """A16 [M,K], B16 [K,N] uint16 fp16 bit patterns; returns C (uint16 bits, or float32)."""
    M, K = A16.shape
    _, N = B16.shape
    if C is None:
        out = np.zeros((M, N), np.float32 if c_f32 else np.uint16)
    else:
        out = np.array(C, copy=True, order="C")
    lib().orc_hgemm_f16(M, N, K, alpha, np.ascontiguousarray(A16), K, np.ascontiguousarray(B16), N,
                        beta, out.ctypes.data, N, int(c_f32), threads)
    return out


def linear_s8(A, Bt, threads=1):
    M, K = A.shape
    N, _ = Bt.shape
    out = np.empty((M, N), np.int32)
    lib().orc_linear_s8(M, N, K, np.ascontiguousarray(A), np.ascontiguousarray(Bt), out, threads)
    return out


def linear_bf16(A16, Bt16, c_f32=False, threads=1):
    M, K = A16.shape
    N, _ = Bt16.shape
    out = np.empty((M, N), np.float32 if c_f32 else np.uint16)
    lib().orc_linear_bf16(M, N, K, np.ascontiguousarray(A16), np.ascontiguousarray(Bt16),
                          out.ctypes.data, int(c_f32), threads)
    return out


def _scalar(fn, *arrs):
    out = np.zeros(1, np.float32)
    fn(arrs[0].size, *arrs, out)
    return out[0]


def sum_f32(x):
    return _scalar(lib().orc_sum_f32, np.ascontiguousarray(x, np.float32))


def dot_f32(a, b):
    return _scalar(lib().orc_dot_f32, np.ascontiguousarray(a, np.float32),
                   np.ascontiguousarray(b, np.float32))


def nrm2_f32(x):
    return _scalar(lib().orc_nrm2_f32, np.ascontiguousarray(x, np.float32))


def dot_f64(a, b=None):
    a = np.ascontiguousarray(a, np.float32)
    r, s = C.c_double(), C.c_double()
    bp = None
    if b is not None:
        b = np.ascontiguousarray(b, np.float32)
        bp = b.ctypes.data
    lib().orc_dot_f64(a.size, a, bp, C.byref(r), C.byref(s))
    return r.value, s.value


def scan_i32(x, seed=0):
    x = np.ascontiguousarray(x, np.int32)
    out = np.empty_like(x)
    lib().orc_scan_i32(x.size, x, out, seed)
    return out


def sum_i32(x):
    x = np.ascontiguousarray(x, np.int32)
    out = np.zeros(1, np.int32)
    lib().orc_sum_i32(x.size, x, out)
    return out[0]


def scan_f32(x):
    x = np.ascontiguousarray(x, np.float32)
    out = np.empty_like(x)
    lib().orc_scan_f32(x.size, x, out)
    return out


def cux_check(ref_arr, x, scale):
    """cux ResultChecker: returns (diff, passed)."""
    r = np.ascontiguousarray(ref_arr, np.float32).reshape(-1)
    v = np.ascontiguousarray(x, np.float32).reshape(-1)
    d = lib().orc_cux_check_f32(r, v, r.size, scale)
    return d, d < 1e-3
