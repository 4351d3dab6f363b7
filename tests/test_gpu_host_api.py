"""GPU: the host-buffer entry points (what a cux plugin given host arrays calls), conversion and
the occupancy hook."""
import ctypes

import numpy as np
import pytest
import torch

import oracle
from hpc_b200 import capi
from hpc_b200 import datagen as dg
from gpu_util import ctx, dev, host, sync

pytestmark = pytest.mark.gpu


def p(a):
    return a.ctypes.data


def test_dot_sum_nrm2_scan_host_buffers():
    c = ctx()
    n = 1_000_003
    a, b = dg.int_fill(n, -2, 2, tag=1), dg.int_fill(n, -2, 2, tag=2)
    r = np.zeros(1, np.float32)
    c.dot_f32_host(n, p(a), p(b), p(r))
    assert r[0] == oracle.dot_f32(a, b)
    c.sum_f32_host(n, p(a), p(r))
    assert r[0] == oracle.sum_f32(a)
    c.nrm2_f32_host(n, p(a), p(r))
    assert r[0] == oracle.nrm2_f32(a)
    x = dg.int_fill(n, -2, 2, np.int32, tag=3)
    out = np.empty_like(x)
    c.scan_i32_host(n, p(x), p(out), 4)
    assert np.array_equal(out, oracle.scan_i32(x, 4))


def test_pinned_host_buffers():
    L = capi.lib()
    n = 1 << 20
    hp = ctypes.c_void_p()
    capi.check(L.cuxb_malloc_host(ctypes.byref(hp), n * 8 + 4))
    buf = np.ctypeslib.as_array(ctypes.cast(hp, ctypes.POINTER(ctypes.c_float)), shape=(2 * n + 1,))
    buf[:n] = dg.int_fill(n, -2, 2, tag=4)
    buf[n:2 * n] = dg.int_fill(n, -2, 2, tag=5)
    ctx().dot_f32_host(n, hp.value, hp.value + 4 * n, hp.value + 8 * n)
    assert buf[2 * n] == oracle.dot_f32(buf[:n].copy(), buf[n:2 * n].copy())
    capi.check(L.cuxb_free_host(hp))


def test_gemm_host_buffers():
    c = ctx()
    M, N, K = 256, 384, 128
    A, B, C = dg.int_fill((M, K), -2, 5, tag=6), dg.int_fill((K, N), -2, 5, tag=7), dg.int_fill((M, N), -2, 5, tag=8)
    want = oracle.gemm_f32(A, B, C, 1.0, 3.0)
    out = C.copy()
    c.sgemm_f32_host(M, N, K, 1.0, p(A), K, p(B), N, 3.0, p(out), N)
    assert np.array_equal(out, want)
    A16, B16, C16 = oracle.f32_to_f16(A), oracle.f32_to_f16(B), oracle.f32_to_f16(C)
    want16 = oracle.hgemm_f16(A16, B16, C16, 1.0, 3.0)
    out16 = C16.copy()
    c.hgemm_f16_host(M, N, K, 1.0, p(A16), K, p(B16), N, 3.0, p(out16), N)
    assert np.array_equal(out16, want16)
    A8, B8 = dg.int_fill((M, K), -128, 127, np.int8, tag=9), dg.int_fill((N, K), -128, 127, np.int8, tag=10)
    out32 = np.empty((M, N), np.int32)
    c.linear_s8_host(M, N, K, p(A8), p(B8), p(out32))
    assert np.array_equal(out32, oracle.linear_s8(A8, B8))


def test_device_convert_matches_rne_oracle():
    c = ctx()
    x = np.concatenate([dg.normal_fill(100000, tag=11) * s for s in (1e-7, 1e-3, 1.0, 3e4, 1e5)])
    d = dev(x)
    h = torch.empty(x.size, dtype=torch.int16, device="cuda")
    c.convert(x.size, d.data_ptr(), capi.FLOAT32, h.data_ptr(), capi.FLOAT16)
    sync()
    assert np.array_equal(host(h, np.uint16), oracle.f32_to_f16(x))
    c.convert(x.size, d.data_ptr(), capi.FLOAT32, h.data_ptr(), capi.BFLOAT16)
    sync()
    assert np.array_equal(host(h, np.uint16), oracle.f32_to_bf16(x))
    back = torch.empty(x.size, dtype=torch.float32, device="cuda")
    c.convert(x.size, h.data_ptr(), capi.BFLOAT16, back.data_ptr(), capi.FLOAT32)
    sync()
    assert np.array_equal(host(back), oracle.bf16_to_f32(oracle.f32_to_bf16(x)))


def test_occupancy_hook_and_launch_counter():
    c = ctx()
    for k, th in (("dot_f32", 512), ("sum_f32", 512), ("nrm2_f32", 512), ("scan_i32", 544), ("sgemm_f32", 256),
                  ("hgemm_f16", 256)):
        ab, occ, threads, smem = c.occupancy(k)
        assert ab >= 1 and 0 < occ <= 1 and threads == th, (k, ab, occ)
    before = capi.launch_count()
    out = torch.zeros(1, device="cuda")
    x = torch.ones(1000, device="cuda")
    c.sum_f32(1000, x.data_ptr(), out.data_ptr())
    sync()
    assert capi.launch_count() == before + 1 and float(out[0]) == 1000.0
    with pytest.raises(capi.CuxbError):
        c.occupancy("nope")


def _splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15))
    x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


def test_device_random_fill_is_a_pure_function_of_seed_and_index():
    c = ctx()
    n, lo, hi, seed = 1_000_003, -2, 5, 12345
    d = torch.empty(n, dtype=torch.float32, device="cuda")
    c.fill_random(n, d.data_ptr(), capi.FLOAT32, lo, hi, 0, seed)
    sync()
    got = host(d)
    with np.errstate(over="ignore"):
        h = _splitmix64(np.uint64(seed) + np.arange(n, dtype=np.uint64) * np.uint64(0xD1342543DE82EF95))
    want = (lo + ((h >> np.uint64(32)) % np.uint64(hi - lo + 1)).astype(np.int64)).astype(np.float32)
    assert np.array_equal(got, want)
    assert got.min() == lo and got.max() == hi and set(np.unique(got)) == set(range(lo, hi + 1))
    # other element types get the same integers; decimals stay inside [lo, hi)
    for dt, tdt in ((capi.INT32, torch.int32), (capi.INT8, torch.int8), (capi.FLOAT16, torch.float16)):
        t = torch.empty(n, dtype=tdt, device="cuda")
        c.fill_random(n, t.data_ptr(), dt, lo, hi, 0, seed)
        sync()
        assert np.array_equal(t.float().cpu().numpy(), want)
    c.fill_random(n, d.data_ptr(), capi.FLOAT32, 0, 3, 2, seed)
    sync()
    g = host(d)
    assert g.min() >= 0 and g.max() < 3 and np.allclose(g * 100, np.round(g * 100), atol=1e-3)


def test_host_pipelines_cut_into_chunks_match_the_device_kernels():
    """Above 32 MiB of traffic the GEMM / scan host entry points become three-stream pipelines (upload of
    chunk i+1, kernel of chunk i, download of chunk i-1).  Same bits as one kernel over device-resident
    data, with padded leading dimensions, beta != 0 (C travels both ways) and a ragged last chunk."""
    c = ctx()
    # ---- fp32, beta != 0, padded rows
    M, N, K = 4224, 1024, 1024   # 33 tile-rows: the last chunk is smaller
    lda, ldb, ldc = K + 8, N + 4, N + 12
    A = np.zeros((M, lda), np.float32); A[:, :K] = dg.int_fill((M, K), -2, 5, tag=41)
    B = np.zeros((K, ldb), np.float32); B[:, :N] = dg.int_fill((K, N), -2, 5, tag=42)
    C = np.full((M, ldc), 9.0, np.float32); C[:, :N] = dg.int_fill((M, N), -2, 5, tag=43)
    dA, dB, dC = dev(A[:, :K].copy()), dev(B[:, :N].copy()), dev(C[:, :N].copy())
    c.sgemm_f32(M, N, K, 2.0, dA.data_ptr(), K, dB.data_ptr(), N, -1.0, dC.data_ptr(), N)
    sync()
    out = C.copy()
    c.sgemm_f32_host(M, N, K, 2.0, p(A), lda, p(B), ldb, -1.0, p(out), ldc)
    assert np.array_equal(out[:, :N], host(dC))
    assert np.all(out[:, N:] == 9.0)                      # the padding of C is left alone
    # ---- fp16 -> fp16
    M, N, K = 8192, 2048, 1024
    A16 = oracle.f32_to_f16(dg.int_fill((M, K), -10, 9, tag=44))
    B16 = oracle.f32_to_f16(dg.int_fill((K, N), -10, 9, tag=45))
    dA, dB = dev(A16), dev(B16)
    dC = torch.zeros((M, N), dtype=torch.int16, device="cuda")
    c.hgemm_f16(M, N, K, 1.0, dA.data_ptr(), K, dB.data_ptr(), N, 0.0, dC.data_ptr(), N, capi.FLOAT16)
    sync()
    out16 = np.zeros((M, N), np.uint16)
    c.hgemm_f16_host(M, N, K, 1.0, p(A16), K, p(B16), N, 0.0, p(out16), N, capi.FLOAT16)
    assert np.array_equal(out16, host(dC, np.uint16))
    # ---- int8 -> int32 (bit-exact against the oracle on sampled rows as well)
    M, N, K = 4352, 2048, 1024
    A8 = dg.int_fill((M, K), -128, 127, tag=46).astype(np.int8)
    Bt = dg.int_fill((N, K), -128, 127, tag=47).astype(np.int8)
    out32 = np.zeros((M, N), np.int32)
    c.linear_s8_host(M, N, K, p(A8), p(Bt), p(out32))
    rows = [0, 1, 255, 256, 2047, 2048, 4095, 4351]
    assert np.array_equal(out32[rows], oracle.linear_s8(A8[rows], Bt, threads=8))
    dA, dB = dev(A8), dev(Bt)
    dC = torch.zeros((M, N), dtype=torch.int32, device="cuda")
    c.linear_s8(M, N, K, dA.data_ptr(), dB.data_ptr(), dC.data_ptr())
    sync()
    assert np.array_equal(out32, host(dC))
    # ---- int32 scan: chunk i continues from chunk i-1's last output, wrap-around included
    n = 6_000_011
    x = np.full(n, 3000, np.int32)
    x[::7] = -11
    out = np.empty_like(x)
    c.scan_i32_host(n, p(x), p(out), -5)
    want = oracle.scan_i32(x, -5)
    assert want.min() < 0 and np.array_equal(out, want)


def test_probes_report_plausible_ceilings():
    """cuxb_probe_hbm / cuxb_probe_fp32 (pocket-ai/eval/cuda/bandwidth_eval.hpp:257-300, max_flops.cuh:11-94 on a
    B200): sanity ranges only -- the numbers themselves go into bench.py's roofline.measured_here."""
    c = ctx()
    rd, wr, cp = c.probe_hbm("read", 1 << 30, 3), c.probe_hbm("write", 1 << 30, 3), c.probe_hbm("copy", 1 << 30, 3)
    assert 4000 < rd < 8200 and 4000 < wr < 8200 and 4000 < cp < 8200, (rd, wr, cp)
    f1, f2 = c.probe_fp32(False, 3), c.probe_fp32(True, 3)
    assert 40 < f1 < 80 and 40 < f2 < 80 and f2 >= 0.95 * f1, (f1, f2)
    with pytest.raises(capi.CuxbError):
        c.probe_hbm("read", 1000, 3)          # too small / not a multiple of 16
    with pytest.raises(capi.CuxbError):
        c.probe_hbm("read", 1 << 20, 3)       # below 64 MiB the loop would barely run: rejected, not mis-reported


def test_tensor_probe_reports_plausible_ceilings():
    """cuxb_probe_tensor: tcgen05.mma on shared-memory-resident operands (the tensor-pipe counterpart of
    max_flops.cuh:11-94).  Dense fp16 is 2250 TFLOP/s nominal, int8 twice that; a power-capped board sits below."""
    c = ctx()
    b16, m16 = c.probe_tensor(capi.FLOAT16, data=1, reps=10)
    bz, _ = c.probe_tensor(capi.BFLOAT16, data=0, reps=5)
    b8, m8 = c.probe_tensor(capi.INT8, data=1, reps=10)
    # (the mean of a few launches can exceed the best of the two single launches before it: clocks ramp up)
    assert 1000 < m16 < 2400 and 1000 < b16 < 2400, (b16, m16)
    assert 1000 < bz < 2400, bz
    assert 2000 < m8 < 4800 and 2000 < b8 < 4800, (b8, m8)
    with pytest.raises(capi.CuxbError):
        c.probe_tensor(capi.FLOAT32)


def test_nvlink_probe_pair():
    """cuxb_probe_nvlink between GPU 0 and GPU 1 (needs two GPUs): every mode lands between PCIe-class and the
    900 GB/s one direction of NVLink 5 can carry."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    c = ctx()
    for mode, depth in (("pull", 0), ("push", 0), ("memcpy", 0), ("bulk", 2), ("bulk", 8)):
        v = c.probe_nvlink(1, mode, depth, 256 << 20, 3)
        assert 100 < v < 950, (mode, depth, v)
    with pytest.raises(capi.CuxbError):
        c.probe_nvlink(0, "pull")             # the context's own GPU is not a peer
