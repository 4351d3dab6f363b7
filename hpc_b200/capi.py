"""ctypes binding of libcuxb.so (include/cuxb.h).

Python is only the test/bench driver: this module adds nothing but argument
marshalling.  There is no fallback -- if the CUDA library is missing or a call
fails, a CuxbError is raised.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# CUXB_LIB_PATH selects a variant build of the same library (tuning experiments only)
LIB_PATH = os.environ.get("CUXB_LIB_PATH") or os.path.join(_HERE, "libcuxb.so")

FLOAT32, INT32, FLOAT16, INT8, BFLOAT16 = 0, 1, 2, 3, 4
MG_EXCHANGE_BYTES = 256       # CUXB_MG_EXCHANGE_BYTES
MG_CHAIN_FLAGS_BYTES = 16640  # CUXB_MG_CHAIN_FLAGS_BYTES
MG_MCAST_FLAGS_BYTES = 16640  # CUXB_MG_MCAST_FLAGS_BYTES


class CuxbError(RuntimeError):
    pass


_vp, _ll, _i, _f, _sz = C.c_void_p, C.c_longlong, C.c_int, C.c_float, C.c_size_t

# name -> (argtypes, restype).  Every symbol declared in include/cuxb.h is listed here;
# tests/test_abi.py checks the header against this table and the .so.
SIGNATURES = {
    "cuxb_ctx_create": ([_i, C.POINTER(_vp)], _i),
    "cuxb_ctx_destroy": ([_vp], _i),
    "cuxb_ctx_device": ([_vp], _i),
    "cuxb_ctx_sm_count": ([_vp], _i),
    "cuxb_last_error_string": ([], C.c_char_p),
    "cuxb_version": ([], C.c_char_p),
    "cuxb_launch_count": ([], _ll),
    "cuxb_malloc": ([C.POINTER(_vp), _sz], _i),
    "cuxb_free": ([_vp], _i),
    "cuxb_malloc_host": ([C.POINTER(_vp), _sz], _i),
    "cuxb_free_host": ([_vp], _i),
    "cuxb_memcpy_h2d": ([_vp, _vp, _sz, _vp], _i),
    "cuxb_memcpy_d2h": ([_vp, _vp, _sz, _vp], _i),
    "cuxb_memcpy_d2d": ([_vp, _vp, _sz, _vp], _i),
    "cuxb_memset": ([_vp, _i, _sz, _vp], _i),
    "cuxb_set_device": ([_i], _i),
    "cuxb_device_count": ([C.POINTER(_i)], _i),
    "cuxb_device_sync": ([], _i),
    "cuxb_device_name": ([_i, C.c_char_p, _sz], _i),
    "cuxb_device_info": ([_i] + [C.POINTER(_i)] * 6 + [C.POINTER(_sz)] * 2, _i),
    "cuxb_stream_create": ([C.POINTER(_vp)], _i),
    "cuxb_stream_destroy": ([_vp], _i),
    "cuxb_stream_sync": ([_vp], _i),
    "cuxb_event_create": ([C.POINTER(_vp)], _i),
    "cuxb_event_destroy": ([_vp], _i),
    "cuxb_event_record": ([_vp, _vp], _i),
    "cuxb_event_sync": ([_vp], _i),
    "cuxb_event_elapsed_ms": ([_vp, _vp, C.POINTER(_f)], _i),
    "cuxb_dot_f32": ([_vp, _vp, _ll, _vp, _vp, _vp], _i),
    "cuxb_sum_f32": ([_vp, _vp, _ll, _vp, _vp], _i),
    "cuxb_nrm2_f32": ([_vp, _vp, _ll, _vp, _vp], _i),
    "cuxb_sum_i32": ([_vp, _vp, _ll, _vp, _vp], _i),
    "cuxb_scan_i32": ([_vp, _vp, _ll, _vp, _vp, C.c_int32], _i),
    "cuxb_scan_f32": ([_vp, _vp, _ll, _vp, _vp, _f], _i),
    "cuxb_scan_i32_noalloc": ([_vp, _ll, _vp, _vp, C.c_int32], _i),
    "cuxb_mg_dot_f32": ([_vp, _vp, _ll, _vp, _vp, _vp, _i, _i, C.POINTER(_vp)], _i),
    "cuxb_mg_sum_f32": ([_vp, _vp, _ll, _vp, _vp, _i, _i, C.POINTER(_vp)], _i),
    "cuxb_mg_nrm2_f32": ([_vp, _vp, _ll, _vp, _vp, _i, _i, C.POINTER(_vp)], _i),
    "cuxb_mg_exscan_total_i32": ([_vp, _vp, _ll, _vp, _vp, _i, _i, C.POINTER(_vp)], _i),
    "cuxb_mg_scan_i32": ([_vp, _vp, _ll, _vp, _vp, _i, _i, C.POINTER(_vp)], _i),
    "cuxb_mg_hgemm_f16_f32acc": ([_vp, _vp, _i, _i, _i, _f, _vp, _i, _vp, _vp, _i, _f, _vp, _i, _i, _i], _i),
    "cuxb_mg_sgemm_f32": ([_vp, _vp, _i, _i, _i, _f, _vp, _i, _vp, _vp, _i, _f, _vp, _i, _i], _i),
    "cuxb_mg_linear_s8_s32": ([_vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _i], _i),
    "cuxb_mg_linear_bf16_f32acc": ([_vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _i, _i], _i),
    "cuxb_mg_linear_f16_f32acc": ([_vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _i, _i], _i),
    "cuxb_mg_set_chain": ([_vp, _vp, _vp, _vp], _i),
    "cuxb_mg_set_multicast": ([_vp, _i, _i, _vp, _vp, _vp], _i),
    "cuxb_enable_peer_access": ([_i, _i], _i),
    "cuxb_mg_create": ([_i, C.POINTER(_i), C.POINTER(_vp)], _i),
    "cuxb_mg_init": ([_i, C.POINTER(_vp)], _i),
    "cuxb_mg_destroy": ([_vp], _i),
    "cuxb_mg_world": ([_vp], _i),
    "cuxb_mg_device": ([_vp, _i], _i),
    "cuxb_mg_ctx": ([_vp, _i], _vp),
    "cuxb_mg_stream": ([_vp, _i], _vp),
    "cuxb_mg_sync": ([_vp], _i),
    "cuxb_mg_all_dot_f32": ([_vp, C.POINTER(_ll), C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp)], _i),
    "cuxb_mg_all_sum_f32": ([_vp, C.POINTER(_ll), C.POINTER(_vp), C.POINTER(_vp)], _i),
    "cuxb_mg_all_nrm2_f32": ([_vp, C.POINTER(_ll), C.POINTER(_vp), C.POINTER(_vp)], _i),
    "cuxb_mg_all_scan_i32": ([_vp, C.POINTER(_ll), C.POINTER(_vp), C.POINTER(_vp)], _i),
    "cuxb_mg_all_sgemm_f32": ([_vp, C.POINTER(_i), _i, _i, _f, C.POINTER(_vp), _i, _vp, _i, _f, C.POINTER(_vp), _i], _i),
    "cuxb_mg_all_hgemm_f16_f32acc": ([_vp, C.POINTER(_i), _i, _i, _f, C.POINTER(_vp), _i, _vp, _i, _f, C.POINTER(_vp), _i, _i], _i),
    "cuxb_mg_all_linear_s8_s32": ([_vp, C.POINTER(_i), _i, _i, C.POINTER(_vp), _vp, C.POINTER(_vp)], _i),
    "cuxb_ipc_export": ([_vp, _vp], _i),
    "cuxb_ipc_import": ([_vp, C.POINTER(_vp)], _i),
    "cuxb_ipc_close": ([_vp], _i),
    "cuxb_sgemm_f32": ([_vp, _vp, _i, _i, _i, _f, _vp, _i, _vp, _i, _f, _vp, _i], _i),
    "cuxb_hgemm_f16_f32acc": ([_vp, _vp, _i, _i, _i, _f, _vp, _i, _vp, _i, _f, _vp, _i, _i], _i),
    "cuxb_gemm_bf16_f32acc": ([_vp, _vp, _i, _i, _i, _f, _vp, _i, _vp, _i, _f, _vp, _i, _i], _i),
    "cuxb_linear_s8_s32": ([_vp, _vp, _i, _i, _i, _vp, _vp, _vp], _i),
    "cuxb_linear_bf16_f32acc": ([_vp, _vp, _i, _i, _i, _vp, _vp, _vp, _i], _i),
    "cuxb_linear_f16_f32acc": ([_vp, _vp, _i, _i, _i, _vp, _vp, _vp, _i], _i),
    "cuxb_convert": ([_vp, _vp, _ll, _vp, _i, _vp, _i], _i),
    "cuxb_fill_random": ([_vp, _vp, _ll, _vp, _i, _i, _i, _i, C.c_ulonglong], _i),
    "cuxb_dot_f32_host": ([_vp, _ll, _vp, _vp, _vp], _i),
    "cuxb_sum_f32_host": ([_vp, _ll, _vp, _vp], _i),
    "cuxb_nrm2_f32_host": ([_vp, _ll, _vp, _vp], _i),
    "cuxb_scan_i32_host": ([_vp, _ll, _vp, _vp, C.c_int32], _i),
    "cuxb_sgemm_f32_host": ([_vp, _i, _i, _i, _f, _vp, _i, _vp, _i, _f, _vp, _i], _i),
    "cuxb_hgemm_f16_f32acc_host": ([_vp, _i, _i, _i, _f, _vp, _i, _vp, _i, _f, _vp, _i, _i], _i),
    "cuxb_linear_s8_s32_host": ([_vp, _i, _i, _i, _vp, _vp, _vp], _i),
    "cuxb_probe_hbm": ([_vp, _i, C.c_size_t, _i, C.POINTER(C.c_double)], _i),
    "cuxb_probe_fp32": ([_vp, _i, _i, C.POINTER(C.c_double)], _i),
    "cuxb_probe_tensor": ([_vp, _i, _i, _i, C.POINTER(C.c_double), C.POINTER(C.c_double)], _i),
    "cuxb_probe_nvlink": ([_vp, _i, _i, _i, C.c_size_t, _i, C.POINTER(C.c_double)], _i),
    "cuxb_workspace_bytes": ([_vp, C.c_char_p, _ll, _ll, _ll, C.POINTER(C.c_size_t)], _i),
    "cuxb_kernel_occupancy": ([_vp, C.c_char_p, C.POINTER(_i), C.POINTER(C.c_double), C.POINTER(_i),
                               C.POINTER(_i)], _i),
}

_lib = None


def lib():
    """Load libcuxb.so and attach prototypes.  Raises if the library is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CuxbError("%s is missing: run `python -m hpc_b200.build` (there is no CPU fallback)" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    for name, (args, res) in SIGNATURES.items():
        fn = getattr(L, name)
        fn.argtypes = args
        fn.restype = res
    L.cuxb_debug_ctx_set.argtypes = [_vp, _i, _i]
    L.cuxb_debug_ctx_set.restype = _i
    _lib = L
    return L


def check(rc, what=""):
    if rc != 0:
        raise CuxbError("%s failed with code %d: %s" % (what or "cuxb call", rc,
                                                        lib().cuxb_last_error_string().decode()))


class Context:
    """cuxb_ctx wrapper: per-device workspace; one per device (and per concurrent stream)."""

    def __init__(self, device=0):
        self._h = _vp()
        check(lib().cuxb_ctx_create(device, C.byref(self._h)), "cuxb_ctx_create")
        self.device = device

    @property
    def handle(self):
        return self._h

    @property
    def sm_count(self):
        return lib().cuxb_ctx_sm_count(self._h)

    def debug_set(self, what, value):
        """Tuning / test hook of THIS context: what = "cta_group" | "group_m" | "wave_sync"."""
        check(lib().cuxb_debug_ctx_set(self._h, {"cta_group": 0, "group_m": 1, "wave_sync": 2, "mc_group": 3, "copy_bufs": 4, "wide_tiles": 5, "direct_epilogue": 6}[what], int(value)),
              "cuxb_debug_ctx_set")

    def close(self):
        if self._h:
            lib().cuxb_ctx_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # --- device-pointer entry points (ptr arguments are ints: tensor.data_ptr()) ---
    def dot_f32(self, n, a, b, res, stream=0):
        check(lib().cuxb_dot_f32(self._h, stream, n, a, b, res), "cuxb_dot_f32")

    def sum_f32(self, n, x, res, stream=0):
        check(lib().cuxb_sum_f32(self._h, stream, n, x, res), "cuxb_sum_f32")

    def nrm2_f32(self, n, x, res, stream=0):
        check(lib().cuxb_nrm2_f32(self._h, stream, n, x, res), "cuxb_nrm2_f32")

    def sum_i32(self, n, x, res, stream=0):
        check(lib().cuxb_sum_i32(self._h, stream, n, x, res), "cuxb_sum_i32")

    def scan_i32(self, n, src, dst, seed=0, stream=0):
        check(lib().cuxb_scan_i32(self._h, stream, n, src, dst, seed), "cuxb_scan_i32")

    def scan_f32(self, n, src, dst, seed=0.0, stream=0):
        check(lib().cuxb_scan_f32(self._h, stream, n, src, dst, seed), "cuxb_scan_f32")

    def sgemm_f32(self, M, N, K, alpha, A, lda, B, ldb, beta, Cp, ldc, stream=0):
        check(lib().cuxb_sgemm_f32(self._h, stream, M, N, K, alpha, A, lda, B, ldb, beta, Cp, ldc),
              "cuxb_sgemm_f32")

    def hgemm_f16(self, M, N, K, alpha, A, lda, B, ldb, beta, Cp, ldc, c_dtype=FLOAT16, stream=0):
        check(lib().cuxb_hgemm_f16_f32acc(self._h, stream, M, N, K, alpha, A, lda, B, ldb, beta, Cp, ldc,
                                          c_dtype), "cuxb_hgemm_f16_f32acc")

    def gemm_bf16(self, M, N, K, alpha, A, lda, B, ldb, beta, Cp, ldc, c_dtype=BFLOAT16, stream=0):
        check(lib().cuxb_gemm_bf16_f32acc(self._h, stream, M, N, K, alpha, A, lda, B, ldb, beta, Cp, ldc,
                                          c_dtype), "cuxb_gemm_bf16_f32acc")

    def linear_s8(self, M, N, K, A, Bt, Cp, stream=0):
        check(lib().cuxb_linear_s8_s32(self._h, stream, M, N, K, A, Bt, Cp), "cuxb_linear_s8_s32")

    def linear_bf16(self, M, N, K, A, Bt, Cp, c_dtype=BFLOAT16, stream=0):
        check(lib().cuxb_linear_bf16_f32acc(self._h, stream, M, N, K, A, Bt, Cp, c_dtype),
              "cuxb_linear_bf16_f32acc")

    def linear_f16(self, M, N, K, A, Bt, Cp, c_dtype=FLOAT16, stream=0):
        check(lib().cuxb_linear_f16_f32acc(self._h, stream, M, N, K, A, Bt, Cp, c_dtype),
              "cuxb_linear_f16_f32acc")

    # --- multi-GPU: reductions / scan with the scalar exchange fused into the kernel (peer memory) ---
    @staticmethod
    def _peers(ptrs):
        return (_vp * len(ptrs))(*ptrs)

    def mg_dot_f32(self, n, a, b, res, world, rank, peer_ptrs, stream=0):
        check(lib().cuxb_mg_dot_f32(self._h, stream, n, a, b, res, world, rank, self._peers(peer_ptrs)), "cuxb_mg_dot_f32")

    def mg_sum_f32(self, n, x, res, world, rank, peer_ptrs, stream=0):
        check(lib().cuxb_mg_sum_f32(self._h, stream, n, x, res, world, rank, self._peers(peer_ptrs)), "cuxb_mg_sum_f32")

    def mg_nrm2_f32(self, n, x, res, world, rank, peer_ptrs, stream=0):
        check(lib().cuxb_mg_nrm2_f32(self._h, stream, n, x, res, world, rank, self._peers(peer_ptrs)), "cuxb_mg_nrm2_f32")

    def mg_scan_i32(self, n, src, dst, world, rank, peer_ptrs, stream=0):
        check(lib().cuxb_mg_scan_i32(self._h, stream, n, src, dst, world, rank, self._peers(peer_ptrs)), "cuxb_mg_scan_i32")

    def mg_hgemm_f16(self, M, N, K, alpha, A, lda, b_root, b_local, ldb, beta, Cp, ldc, is_root, c_dtype=FLOAT16,
                     stream=0):
        check(lib().cuxb_mg_hgemm_f16_f32acc(self._h, stream, M, N, K, alpha, A, lda, b_root, b_local, ldb, beta, Cp,
                                             ldc, c_dtype, int(is_root)), "cuxb_mg_hgemm_f16_f32acc")

    def mg_sgemm_f32(self, M, N, K, alpha, A, lda, b_root, b_local, ldb, beta, Cp, ldc, is_root, stream=0):
        check(lib().cuxb_mg_sgemm_f32(self._h, stream, M, N, K, alpha, A, lda, b_root, b_local, ldb, beta, Cp, ldc,
                                      int(is_root)), "cuxb_mg_sgemm_f32")

    def mg_linear_s8(self, M, N, K, A, bt_root, bt_local, Cp, is_root, stream=0):
        check(lib().cuxb_mg_linear_s8_s32(self._h, stream, M, N, K, A, bt_root, bt_local, Cp, int(is_root)),
              "cuxb_mg_linear_s8_s32")

    def mg_linear_16(self, kind, M, N, K, A, bt_root, bt_local, Cp, c_dtype, is_root, stream=0):
        f = {"bf16": lib().cuxb_mg_linear_bf16_f32acc, "f16": lib().cuxb_mg_linear_f16_f32acc}[kind]
        check(f(self._h, stream, M, N, K, A, bt_root, bt_local, Cp, c_dtype, int(is_root)), "cuxb_mg_linear_%s_f32acc" % kind)

    def mg_set_chain(self, flags_local, flags_up, flags_down):
        check(lib().cuxb_mg_set_chain(self._h, flags_local or None, flags_up or None, flags_down or None),
              "cuxb_mg_set_chain")

    def mg_set_multicast(self, world, rank, flags_local, flags_mc, b_mc):
        check(lib().cuxb_mg_set_multicast(self._h, world, rank, flags_local or None, flags_mc or None, b_mc or None),
              "cuxb_mg_set_multicast")

    def convert(self, n, src, src_dtype, dst, dst_dtype, stream=0):
        check(lib().cuxb_convert(self._h, stream, n, src, src_dtype, dst, dst_dtype), "cuxb_convert")

    def fill_random(self, n, dst, dtype, lo, hi, decimal_pose=0, seed=20240607, stream=0):
        check(lib().cuxb_fill_random(self._h, stream, n, dst, dtype, lo, hi, decimal_pose, seed), "cuxb_fill_random")

    def probe_hbm(self, which, nbytes=2 << 30, reps=5):
        """Measured HBM ceiling in GB/s: which = "read" | "write" | "copy"."""
        v = C.c_double()
        check(lib().cuxb_probe_hbm(self._h, {"read": 0, "write": 1, "copy": 2}[which], nbytes, reps, C.byref(v)),
              "cuxb_probe_hbm")
        return v.value

    def probe_fp32(self, packed=True, reps=5):
        """Measured FP32 rate of the CUDA cores in TFLOP/s (scalar FFMA or packed FFMA2)."""
        v = C.c_double()
        check(lib().cuxb_probe_fp32(self._h, int(packed), reps, C.byref(v)), "cuxb_probe_fp32")
        return v.value

    def probe_tensor(self, kind=FLOAT16, data=1, reps=20):
        """Measured tensor-pipe ceiling (TFLOP/s, TOP/s for INT8): (burst, mean over `reps` back-to-back launches)."""
        b, m = C.c_double(), C.c_double()
        check(lib().cuxb_probe_tensor(self._h, kind, int(data), reps, C.byref(b), C.byref(m)), "cuxb_probe_tensor")
        return b.value, m.value

    def probe_nvlink(self, peer_device, mode="bulk", depth=2, nbytes=256 << 20, reps=5):
        """GB/s between peer_device's memory and this GPU: mode = "pull" | "push" | "memcpy" | "bulk"."""
        v = C.c_double()
        check(lib().cuxb_probe_nvlink(self._h, peer_device, {"pull": 0, "push": 1, "memcpy": 2, "bulk": 3}[mode], depth,
                                      nbytes, reps, C.byref(v)), "cuxb_probe_nvlink")
        return v.value

    def occupancy(self, kernel):
        ab, occ, th, sm = _i(), C.c_double(), _i(), _i()
        check(lib().cuxb_kernel_occupancy(self._h, kernel.encode(), C.byref(ab), C.byref(occ), C.byref(th),
                                          C.byref(sm)), "cuxb_kernel_occupancy")
        return ab.value, occ.value, th.value, sm.value

    # --- host-buffer entry points (numpy arrays or raw host pointers) ---
    def dot_f32_host(self, n, a_ptr, b_ptr, res_ptr):
        check(lib().cuxb_dot_f32_host(self._h, n, a_ptr, b_ptr, res_ptr), "cuxb_dot_f32_host")

    def sum_f32_host(self, n, x_ptr, res_ptr):
        check(lib().cuxb_sum_f32_host(self._h, n, x_ptr, res_ptr), "cuxb_sum_f32_host")

    def nrm2_f32_host(self, n, x_ptr, res_ptr):
        check(lib().cuxb_nrm2_f32_host(self._h, n, x_ptr, res_ptr), "cuxb_nrm2_f32_host")

    def scan_i32_host(self, n, in_ptr, out_ptr, seed=0):
        check(lib().cuxb_scan_i32_host(self._h, n, in_ptr, out_ptr, seed), "cuxb_scan_i32_host")

    def sgemm_f32_host(self, M, N, K, alpha, A, lda, B, ldb, beta, Cp, ldc):
        check(lib().cuxb_sgemm_f32_host(self._h, M, N, K, alpha, A, lda, B, ldb, beta, Cp, ldc),
              "cuxb_sgemm_f32_host")

    def hgemm_f16_host(self, M, N, K, alpha, A, lda, B, ldb, beta, Cp, ldc, c_dtype=FLOAT16):
        check(lib().cuxb_hgemm_f16_f32acc_host(self._h, M, N, K, alpha, A, lda, B, ldb, beta, Cp, ldc, c_dtype),
              "cuxb_hgemm_f16_f32acc_host")

    def linear_s8_host(self, M, N, K, A, Bt, Cp):
        check(lib().cuxb_linear_s8_s32_host(self._h, M, N, K, A, Bt, Cp), "cuxb_linear_s8_s32_host")


def launch_count():
    return lib().cuxb_launch_count()


class MultiGpu:
    """cuxb_mg wrapper: one host thread driving `devices` (see include/cuxb.h).  Arguments of the all_* calls
    are lists indexed by rank: element counts / row counts and device pointers (ints) on that rank's GPU."""

    def __init__(self, devices):
        devices = list(devices)
        self._h = _vp()
        arr = (_i * len(devices))(*devices)
        check(lib().cuxb_mg_create(len(devices), arr, C.byref(self._h)), "cuxb_mg_create")
        self.world = len(devices)
        self.devices = devices

    def close(self):
        if self._h:
            lib().cuxb_mg_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def stream(self, rank):
        return lib().cuxb_mg_stream(self._h, rank)

    def sync(self):
        check(lib().cuxb_mg_sync(self._h), "cuxb_mg_sync")

    def _ptrs(self, ps):
        assert len(ps) == self.world
        return (_vp * self.world)(*ps)

    def _lens(self, ns):
        return (_ll * self.world)(*ns)

    def _rows(self, ms):
        return (_i * self.world)(*ms)

    def all_dot_f32(self, n, a, b, res):
        check(lib().cuxb_mg_all_dot_f32(self._h, self._lens(n), self._ptrs(a), self._ptrs(b), self._ptrs(res)), "cuxb_mg_all_dot_f32")

    def all_sum_f32(self, n, x, res):
        check(lib().cuxb_mg_all_sum_f32(self._h, self._lens(n), self._ptrs(x), self._ptrs(res)), "cuxb_mg_all_sum_f32")

    def all_nrm2_f32(self, n, x, res):
        check(lib().cuxb_mg_all_nrm2_f32(self._h, self._lens(n), self._ptrs(x), self._ptrs(res)), "cuxb_mg_all_nrm2_f32")

    def all_scan_i32(self, n, src, dst):
        check(lib().cuxb_mg_all_scan_i32(self._h, self._lens(n), self._ptrs(src), self._ptrs(dst)), "cuxb_mg_all_scan_i32")

    def all_sgemm_f32(self, M, N, K, alpha, A, lda, B_root, ldb, beta, Cp, ldc):
        check(lib().cuxb_mg_all_sgemm_f32(self._h, self._rows(M), N, K, alpha, self._ptrs(A), lda, B_root, ldb, beta,
                                          self._ptrs(Cp), ldc), "cuxb_mg_all_sgemm_f32")

    def all_hgemm_f16(self, M, N, K, alpha, A, lda, B_root, ldb, beta, Cp, ldc, c_dtype=FLOAT16):
        check(lib().cuxb_mg_all_hgemm_f16_f32acc(self._h, self._rows(M), N, K, alpha, self._ptrs(A), lda, B_root, ldb, beta,
                                                 self._ptrs(Cp), ldc, c_dtype), "cuxb_mg_all_hgemm_f16_f32acc")

    def all_linear_s8(self, M, N, K, A, Bt_root, Cp):
        check(lib().cuxb_mg_all_linear_s8_s32(self._h, self._rows(M), N, K, self._ptrs(A), Bt_root, self._ptrs(Cp)),
              "cuxb_mg_all_linear_s8_s32")


def ipc_export(dptr):
    """64-byte handle (bytes) of a cuxb_malloc'ed device buffer, for another process to cuxb_ipc_import."""
    buf = C.create_string_buffer(64)
    check(lib().cuxb_ipc_export(dptr, buf), "cuxb_ipc_export")
    return buf.raw


def ipc_import(handle):
    out = _vp()
    check(lib().cuxb_ipc_import(C.c_char_p(handle), C.byref(out)), "cuxb_ipc_import")
    return out.value


def dmalloc(nbytes):
    out = _vp()
    check(lib().cuxb_malloc(C.byref(out), nbytes), "cuxb_malloc")
    return out.value
