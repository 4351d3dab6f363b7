"""GPU parity: fp32 SGEMM (CUDA cores) and the tcgen05 tensor-core GEMMs, through the C ABI.

Bar:
  * int8 linear: BIT-EXACT vs LinearNormalInt8 (simd/x86/linear.cpp:199-211).
  * fp32 / fp16 / bf16 with the reference's integer-valued fills: every product and partial sum
    is exactly representable (|partial| < 2^24), so any summation order gives the same bits ->
    BIT-EXACT vs GemmHostV0 (gemm.cpp:13-34) / GemmHostV1-half (gemm_fp16_wmma.cu:75-90) /
    LinearNormalBf16 (linear.cpp:137-151).
  * general float fills: order-of-summation note -- the GPU adds K in tile order with FMA / tensor-
    core accumulation, the reference serially.  Tolerance: |gpu - ref| <= RTOL_K * sqrt(K)-free
    bound  2 * K * eps32 * max|a| * max|b|  for fp32 accumulate, and for 16-bit outputs one extra
    half-ulp of the output format (relative 2^-10 fp16, 2^-7 bf16).  cux's own MAE < 1e-3 rule
    (op_assistor.h:38-46) is checked as well.
"""
import os

import numpy as np
import pytest
import torch

import oracle
from hpc_b200 import capi
from hpc_b200 import datagen as dg
from gpu_util import bits32, ctx, dev, host, sync

pytestmark = pytest.mark.gpu
EPS = float(np.finfo(np.float32).eps)
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_vectors.npz"))


# ------------------------------------------------------------------ fp32 SGEMM
def run_sgemm(A, B, C, alpha, beta):
    M, K = A.shape
    _, N = B.shape
    dA, dB, dC = dev(A), dev(B), dev(C)
    ctx().sgemm_f32(M, N, K, alpha, dA.data_ptr(), K, dB.data_ptr(), N, beta, dC.data_ptr(), N)
    sync()
    return host(dC)


@pytest.mark.parametrize("shape", [(128, 128, 128), (256, 384, 512), (512, 576, 480), (1, 1, 1), (100, 70, 33),
                                   (129, 257, 65), (130, 128, 20), (128, 132, 16)])
@pytest.mark.parametrize("ab", [(1.0, 3.0), (1.0, 0.0), (2.0, -1.0)])
def test_sgemm_integer_fill_bit_exact(shape, ab):
    M, N, K = shape
    A, B, C = dg.int_fill((M, K), -2, 5, tag=1), dg.int_fill((K, N), -2, 5, tag=2), dg.int_fill((M, N), -2, 5, tag=3)
    got = run_sgemm(A, B, C, *ab)
    want = oracle.gemm_f32(A, B, C, *ab, order="ikj", threads=8)
    assert np.array_equal(bits32(got), bits32(want))


@pytest.mark.parametrize("shape", [(256, 256, 256), (200, 136, 1000)])
def test_sgemm_float_fill_tolerance(shape):
    M, N, K = shape
    A, B, C = dg.uniform_fill((M, K), tag=4), dg.uniform_fill((K, N), tag=5), dg.uniform_fill((M, N), tag=6)
    got = run_sgemm(A, B, C, 1.0, 3.0)
    want = oracle.gemm_f32(A, B, C, 1.0, 3.0, order="ikj", threads=8)
    assert np.max(np.abs(got - want)) <= 2 * K * EPS * 1.0 * 1.0 + 4 * EPS * 3.0
    d, ok = oracle.cux_check(want, got, 1.0 / (M * N))
    assert ok, d


@pytest.mark.parametrize("shape", [(128, 128, 1), (130, 70, 1)])   # vectorised and guarded epilogues
def test_sgemm_float_fill_k1_beta_bit_exact(shape):
    # K = 1: there is no summation order to differ in, so with alpha = 1 the result must equal
    # GemmHostV0's fl(fl(beta*c) + fl(a*b)) bit for bit on arbitrary floats -- which it only does if the
    # epilogue rounds alpha*acc and beta*c separately (no fma contraction).
    M, N, K = shape
    A, B, C = dg.uniform_fill((M, K), tag=14), dg.uniform_fill((K, N), tag=15), dg.uniform_fill((M, N), tag=16)
    for beta in (0.75, -3.0, 1.0 / 3.0):
        got = run_sgemm(A, B, C, 1.0, beta)
        want = oracle.gemm_f32(A, B, C, 1.0, beta, order="v0")
        assert np.array_equal(bits32(got), bits32(want)), beta


def test_sgemm_golden_and_leading_dimensions():
    for name in ("gemm_int", "gemm_flt"):
        got = run_sgemm(G[name + "_A"], G[name + "_B"], G[name + "_C0"], 1.0, 3.0)
        if name == "gemm_int":
            assert np.array_equal(bits32(got), bits32(G[name + "_C"]))
        else:
            assert np.allclose(got, G[name + "_C"], rtol=0, atol=2 * 80 * EPS + 16 * EPS)
    # sub-matrix views: lda/ldb/ldc larger than the row length
    M, N, K = 128, 128, 64
    Abig, Bbig, Cbig = dg.int_fill((M, K + 12), -2, 5, tag=7), dg.int_fill((K, N + 4), -2, 5, tag=8), dg.int_fill((M, N + 8), -2, 5, tag=9)
    dA, dB, dC = dev(Abig), dev(Bbig), dev(Cbig)
    ctx().sgemm_f32(M, N, K, 1.0, dA.data_ptr(), K + 12, dB.data_ptr(), N + 4, 3.0, dC.data_ptr(), N + 8)
    sync()
    got = host(dC)
    want = oracle.gemm_f32(Abig[:, :K].copy(), Bbig[:, :N].copy(), Cbig[:, :N].copy(), 1.0, 3.0)
    assert np.array_equal(got[:, :N], want)
    assert np.array_equal(got[:, N:], Cbig[:, N:])  # padding untouched


def test_sgemm_config1_1024_cube():
    # BASELINE config 1: cux gemm fp32 M=N=K=1024, alpha=1 beta=3, integer fill [-2,5] (demo.cpp:77,97)
    n = 1024
    A, B, C = dg.int_fill((n, n), -2, 5, tag=11), dg.int_fill((n, n), -2, 5, tag=12), dg.int_fill((n, n), -2, 5, tag=13)
    got = run_sgemm(A, B, C, 1.0, 3.0)
    want = oracle.gemm_f32(A, B, C, 1.0, 3.0, order="ikj", threads=8)
    assert np.array_equal(bits32(got), bits32(want))


def test_sgemm_full_size_8192_sampled_rows_and_linearity():
    # BASELINE config 3: 8192^3.  Sampled rows vs the oracle (rows are independent) + linearity:
    # gemm(A, B, alpha=2) == 2 * gemm(A, B) exactly for integer data.
    n = 8192
    A, B = dg.int_fill((n, n), -2, 5, tag=21), dg.int_fill((n, n), -2, 5, tag=22)
    dA, dB = dev(A), dev(B)
    dC = torch.zeros((n, n), device="cuda")
    ctx().sgemm_f32(n, n, n, 1.0, dA.data_ptr(), n, dB.data_ptr(), n, 0.0, dC.data_ptr(), n)
    sync()
    rows = [0, 1, 127, 128, 4095, 4096, 8190, 8191]
    want = oracle.gemm_f32(A[rows].copy(), B, np.zeros((len(rows), n), np.float32), 1.0, 0.0, order="ikj", threads=8)
    assert np.array_equal(host(dC[rows]), want)
    dC2 = torch.zeros((n, n), device="cuda")
    ctx().sgemm_f32(n, n, n, 2.0, dA.data_ptr(), n, dB.data_ptr(), n, 0.0, dC2.data_ptr(), n)
    sync()
    assert torch.equal(dC2, 2 * dC)


# ------------------------------------------------------------------ fp16 GEMM (tcgen05)
def run_hgemm(A16, B16, C0=None, alpha=1.0, beta=0.0, c_f32=False):
    M, K = A16.shape
    _, N = B16.shape
    dA, dB = dev(A16), dev(B16)
    if C0 is None:
        dC = torch.zeros((M, N), dtype=torch.float32 if c_f32 else torch.int16, device="cuda")
    else:
        dC = dev(C0)
    ctx().hgemm_f16(M, N, K, alpha, dA.data_ptr(), K, dB.data_ptr(), N, beta, dC.data_ptr(), N,
                    capi.FLOAT32 if c_f32 else capi.FLOAT16)
    sync()
    return host(dC) if c_f32 else host(dC, np.uint16)


@pytest.fixture(params=[2, 1], ids=["cta_pair", "single_cta"])
def cta_group(request):
    ctx().debug_set("cta_group", request.param)   # per context: no other context's kernels change
    yield request.param
    ctx().debug_set("cta_group", 0)


@pytest.mark.parametrize("shape", [(128, 128, 64), (256, 256, 256), (512, 768, 1024), (200, 136, 72), (1, 8, 8),
                                   (300, 520, 200), (1024, 1024, 2048)])
def test_hgemm_integer_fill_bit_exact(shape, cta_group):
    M, N, K = shape
    A16 = oracle.f32_to_f16(dg.int_fill((M, K), -10, 9, tag=31))  # gemm_fp16_wmma.cu:46
    B16 = oracle.f32_to_f16(dg.int_fill((K, N), -10, 9, tag=32))
    want = oracle.hgemm_f16(A16, B16, threads=8)
    got = run_hgemm(A16, B16)
    assert np.array_equal(got, want)
    # fp32 output = the raw fp32 accumulator, exact
    want32 = oracle.hgemm_f16(A16, B16, c_f32=True, threads=8)
    got32 = run_hgemm(A16, B16, c_f32=True)
    assert np.array_equal(bits32(got32), bits32(want32))


def test_hgemm_alpha_beta_cux_convention(cta_group):
    # cux runs gemm with alpha=1, beta=3 on fp16 storage (demo.cpp:77, gemm.cu:266-269)
    M, N, K = 256, 512, 128
    A16 = oracle.f32_to_f16(dg.int_fill((M, K), -2, 5, tag=33))
    B16 = oracle.f32_to_f16(dg.int_fill((K, N), -2, 5, tag=34))
    C16 = oracle.f32_to_f16(dg.int_fill((M, N), -2, 5, tag=35))
    want = oracle.hgemm_f16(A16, B16, C16, 1.0, 3.0, threads=8)
    assert np.array_equal(run_hgemm(A16, B16, C16, 1.0, 3.0), want)
    C32 = dg.int_fill((M, N), -2, 5, tag=36)
    want = oracle.hgemm_f16(A16, B16, C32, 2.0, -1.0, c_f32=True, threads=8)
    assert np.array_equal(run_hgemm(A16, B16, C32, 2.0, -1.0, c_f32=True), want)


def test_hgemm_normal_fill_tolerance_and_golden(cta_group):
    for name in ("hgemm_int", "hgemm_nrm"):
        got = run_hgemm(G[name + "_A16"], G[name + "_B16"])
        if name == "hgemm_int":
            assert np.array_equal(got, G[name + "_C16"])
        else:
            g, w = oracle.f16_to_f32(got), oracle.f16_to_f32(G[name + "_C16"])
            assert np.all(np.abs(g - w) <= 2.0 ** -10 * np.abs(w) + 64 * 2 * EPS * 16)
    M, N, K = 384, 256, 4096
    A16, B16 = oracle.f32_to_f16(dg.normal_fill((M, K), tag=37)), oracle.f32_to_f16(dg.normal_fill((K, N), tag=38))
    want = oracle.hgemm_f16(A16, B16, c_f32=True, threads=8)
    got = run_hgemm(A16, B16, c_f32=True)
    # fp32 accumulate, different order: |diff| <= 2*K*eps*max|a||b| is very loose; check a tight empirical one too
    assert np.max(np.abs(got - want)) <= 1e-5 * np.sqrt(K) * 16
    g16, w16 = oracle.f16_to_f32(run_hgemm(A16, B16)), oracle.f16_to_f32(oracle.hgemm_f16(A16, B16, threads=8))
    assert np.all(np.abs(g16 - w16) <= 2.0 ** -10 * np.abs(w16) + 1e-3)
    d, ok = oracle.cux_check(w16, g16, 1.0 / (M * N))
    assert ok, d


def test_hgemm_full_size_16384_sampled_rows():
    # BASELINE config 4: 16384^3 fp16, fp32 accumulate.  Integer fill [-10,9]; rows are independent,
    # so a sample of rows is compared bit-for-bit with the oracle; a checksum ties in the rest.
    n = 16384
    g = torch.Generator(device="cuda")
    g.manual_seed(dg.SEED)
    dA = torch.randint(-10, 10, (n, n), device="cuda", generator=g, dtype=torch.int32).to(torch.float16)
    dB = torch.randint(-10, 10, (n, n), device="cuda", generator=g, dtype=torch.int32).to(torch.float16)
    dC = torch.zeros((n, n), dtype=torch.float32, device="cuda")
    ctx().hgemm_f16(n, n, n, 1.0, dA.data_ptr(), n, dB.data_ptr(), n, 0.0, dC.data_ptr(), n, capi.FLOAT32)
    sync()
    rows = [0, 127, 128, 255, 256, 8191, 16383]
    A16 = host(dA[rows].view(torch.int16), np.uint16)
    B16 = host(dB.view(torch.int16), np.uint16)
    want = oracle.hgemm_f16(A16, B16, c_f32=True, threads=8)
    assert np.array_equal(host(dC[rows]), want)
    # the same with fp16 OUTPUT, cux's convention (gemm.cu:266-269) and what bench.py times: at K = 16384
    # |acc| goes far beyond 2048, so the final round-to-nearest-even matters on most elements
    dC16 = torch.zeros((n, n), dtype=torch.float16, device="cuda")
    ctx().hgemm_f16(n, n, n, 1.0, dA.data_ptr(), n, dB.data_ptr(), n, 0.0, dC16.data_ptr(), n, capi.FLOAT16)
    sync()
    want16 = oracle.hgemm_f16(A16, B16, threads=8)
    assert np.array_equal(host(dC16[rows].view(torch.int16), np.uint16), want16)
    assert np.abs(want).max() > 2048
    assert torch.equal(dC16, dC.to(torch.float16))   # every element: one RNE of the fp32 accumulator
    # checksum of checksums: sum_j C[i,j] == A[i,:] . rowsum(B)  (exact in fp64 for these integers)
    rs = dB.double().sum(dim=1)
    lhs = dC.double().sum(dim=1)
    rhs = dA.double() @ rs
    assert torch.equal(lhs, rhs)


# ------------------------------------------------------------------ linear (A * Bt^T)
@pytest.mark.parametrize("shape", [(128, 128, 128), (200, 300, 400), (8, 8, 16), (513, 260, 272), (1024, 2048, 1024)])
def test_linear_int8_bit_exact(shape, cta_group):
    M, N, K = shape
    A = dg.int_fill((M, K), -128, 127, np.int8, tag=41)
    B = dg.int_fill((N, K), -128, 127, np.int8, tag=42)
    dA, dB = dev(A), dev(B)
    dC = torch.full((M, N), -1, dtype=torch.int32, device="cuda")
    ctx().linear_s8(M, N, K, dA.data_ptr(), dB.data_ptr(), dC.data_ptr())
    sync()
    assert np.array_equal(host(dC), oracle.linear_s8(A, B, threads=8))


def test_linear_int8_golden_and_sample_fill(cta_group):
    dA, dB = dev(G["lin8_A"]), dev(G["lin8_Bt"])
    dC = torch.zeros((128, 128), dtype=torch.int32, device="cuda")
    ctx().linear_s8(128, 128, 128, dA.data_ptr(), dB.data_ptr(), dC.data_ptr())
    sync()
    assert np.array_equal(host(dC), G["lin8_C"])
    # the sample's own data (i%9 + j%9, linear.cpp:19-25) at its own shape M=200 N=300 K=400
    A, B = dg.linear_sample_fill(200, 400, np.int8), dg.linear_sample_fill(300, 400, np.int8)
    dA, dB = dev(A), dev(B)
    dC = torch.zeros((200, 300), dtype=torch.int32, device="cuda")
    ctx().linear_s8(200, 300, 400, dA.data_ptr(), dB.data_ptr(), dC.data_ptr())
    sync()
    assert np.array_equal(host(dC), oracle.linear_s8(A, B))


def test_linear_int8_full_size_8192_sampled_rows():
    # BASELINE config 5a: M=N=K=8192, A in [0,16] (linear.cpp:22), B in [-8,8]; plus signed A.
    n = 8192
    for tag, lo in ((51, 0), (53, -8)):
        A = dg.int_fill((n, n), lo, 16 if lo == 0 else 8, np.int8, tag=tag)
        B = dg.int_fill((n, n), -8, 8, np.int8, tag=tag + 1)
        dA, dB = dev(A), dev(B)
        dC = torch.zeros((n, n), dtype=torch.int32, device="cuda")
        ctx().linear_s8(n, n, n, dA.data_ptr(), dB.data_ptr(), dC.data_ptr())
        sync()
        rows = [0, 1, 127, 128, 4097, 8191]
        assert np.array_equal(host(dC[rows]), oracle.linear_s8(A[rows].copy(), B, threads=8))
        lhs = dC.sum(dim=1, dtype=torch.int64)
        rhs = (dA.to(torch.float64) @ dB.to(torch.float64).sum(dim=0)).to(torch.int64)
        assert torch.equal(lhs, rhs)


@pytest.mark.parametrize("kind", ["bf16", "f16"])
def test_linear_16bit(kind, cta_group):
    M, N, K = 256, 384, 512
    Af, Bf = dg.int_fill((M, K), -4, 4, tag=61), dg.int_fill((N, K), -4, 4, tag=62)
    if kind == "bf16":
        A16, B16 = oracle.f32_to_bf16(Af), oracle.f32_to_bf16(Bf)
        want = oracle.linear_bf16(A16, B16, threads=8)
    else:
        A16, B16 = oracle.f32_to_f16(Af), oracle.f32_to_f16(Bf)
        want = oracle.hgemm_f16(A16, np.ascontiguousarray(B16.T), threads=8)
    dA, dB = dev(A16), dev(B16)
    dC = torch.zeros((M, N), dtype=torch.int16, device="cuda")
    if kind == "bf16":
        ctx().linear_bf16(M, N, K, dA.data_ptr(), dB.data_ptr(), dC.data_ptr())
    else:
        ctx().linear_f16(M, N, K, dA.data_ptr(), dB.data_ptr(), dC.data_ptr())
    sync()
    assert np.array_equal(host(dC, np.uint16), want)  # integer fill: bit-exact
    if kind == "bf16":
        # golden (normal fill) within one bf16 ulp of the serial reference
        dA, dB = dev(G["linbf_A"]), dev(G["linbf_Bt"])
        dC = torch.zeros((128, 128), dtype=torch.int16, device="cuda")
        ctx().linear_bf16(128, 128, 128, dA.data_ptr(), dB.data_ptr(), dC.data_ptr())
        sync()
        g, w = oracle.bf16_to_f32(host(dC, np.uint16)), oracle.bf16_to_f32(G["linbf_C"])
        assert np.all(np.abs(g - w) <= 2.0 ** -7 * np.abs(w) + 1e-3)


def test_gemm_argument_errors_are_loud():
    c = ctx()
    t = torch.zeros(64, device="cuda")
    with pytest.raises(capi.CuxbError):
        c.hgemm_f16(16, 16, 16, 1.0, t.data_ptr(), 16, t.data_ptr(), 16, 0.0, t.data_ptr(), 12)  # ldc*2 % 16 != 0
    with pytest.raises(capi.CuxbError):
        c.sgemm_f32(16, 16, 16, 1.0, t.data_ptr(), 8, t.data_ptr(), 16, 0.0, t.data_ptr(), 16)  # lda < K
    with pytest.raises(capi.CuxbError):
        c.linear_s8(16, 16, 8, t.data_ptr(), t.data_ptr(), t.data_ptr())  # K % 16 != 0


# ------------------------------------------------------------------ further edge cases
def test_hgemm_submatrix_leading_dimensions(cta_group):
    # A, B, C are windows of larger row-major arrays (lda > K, ldb > N, ldc > N), as a caller slicing
    # a bigger Array4D would pass them
    M, N, K = 200, 264, 136
    lda, ldb, ldc = K + 24, N + 40, N + 8
    Ab = oracle.f32_to_f16(dg.int_fill((M, lda), -10, 9, tag=71))
    Bb = oracle.f32_to_f16(dg.int_fill((K, ldb), -10, 9, tag=72))
    Cb = oracle.f32_to_f16(dg.int_fill((M, ldc), -3, 3, tag=73))
    dA, dB, dC = dev(Ab), dev(Bb), dev(Cb)
    ctx().hgemm_f16(M, N, K, 1.0, dA.data_ptr(), lda, dB.data_ptr(), ldb, 2.0, dC.data_ptr(), ldc, capi.FLOAT16)
    sync()
    got = host(dC, np.uint16)
    want = oracle.hgemm_f16(np.ascontiguousarray(Ab[:, :K]), np.ascontiguousarray(Bb[:, :N]),
                            np.ascontiguousarray(Cb[:, :N]), 1.0, 2.0, threads=8)
    assert np.array_equal(got[:, :N], want)
    assert np.array_equal(got[:, N:], Cb[:, N:])  # columns outside the window untouched


@pytest.mark.parametrize("K", [8, 56, 64, 72, 4104])
def test_hgemm_k_tails(K, cta_group):
    M, N = 130, 264
    A16 = oracle.f32_to_f16(dg.int_fill((M, K), -4, 4, tag=74))
    B16 = oracle.f32_to_f16(dg.int_fill((K, N), -4, 4, tag=75))
    assert np.array_equal(run_hgemm(A16, B16, c_f32=True), oracle.hgemm_f16(A16, B16, c_f32=True, threads=8))


def test_int8_extremes_and_minimal_k(cta_group):
    # -128 * -128 summed over K: the largest magnitudes the accumulator sees
    for (M, N, K), fill in (((128, 256, 8192), -128), ((16, 16, 16), 127), ((3, 4, 16), -128)):
        A = np.full((M, K), fill, np.int8)
        B = np.full((N, K), -128, np.int8)
        B[::2] = 127
        dA, dB = dev(A), dev(B)
        dC = torch.zeros((M, N), dtype=torch.int32, device="cuda")
        ctx().linear_s8(M, N, K, dA.data_ptr(), dB.data_ptr(), dC.data_ptr())
        sync()
        assert np.array_equal(host(dC), oracle.linear_s8(A, B, threads=8))


def test_linear_bf16_normal_fill_large_within_one_ulp(cta_group):
    M, N, K = 512, 512, 2048
    A16 = oracle.f32_to_bf16(dg.normal_fill((M, K), tag=76))
    B16 = oracle.f32_to_bf16(dg.normal_fill((N, K), tag=77))
    dA, dB = dev(A16), dev(B16)
    dC = torch.zeros((M, N), dtype=torch.float32, device="cuda")
    ctx().linear_bf16(M, N, K, dA.data_ptr(), dB.data_ptr(), dC.data_ptr(), capi.FLOAT32)
    sync()
    want = oracle.linear_bf16(A16, B16, c_f32=True, threads=8)
    # bf16 products are exact in fp32; only the order of the fp32 additions differs
    assert np.max(np.abs(host(dC) - want)) <= 1e-5 * np.sqrt(K) * 16


def test_sgemm_alpha_zero_and_beta_one_leave_c():
    M, N, K = 256, 128, 64
    A, B, C = dg.uniform_fill((M, K), tag=78), dg.uniform_fill((K, N), tag=79), dg.uniform_fill((M, N), tag=80)
    assert np.array_equal(run_sgemm(A, B, C, 0.0, 1.0), C)
    assert np.array_equal(run_sgemm(A, B, C, 0.0, 0.0), np.zeros_like(C))


def test_gemm_bf16_cux_layout(cta_group):
    # bf16 operands, B stored [K,N] like cux's gemm: equals LinearNormalBf16 on B^T
    M, N, K = 384, 520, 200
    Af, Bf = dg.int_fill((M, K), -6, 6, tag=81), dg.int_fill((K, N), -6, 6, tag=82)
    A16, B16 = oracle.f32_to_bf16(Af), oracle.f32_to_bf16(Bf)
    want = oracle.linear_bf16(A16, np.ascontiguousarray(B16.T), threads=8)
    dA, dB = dev(A16), dev(B16)
    dC = torch.zeros((M, N), dtype=torch.int16, device="cuda")
    ctx().gemm_bf16(M, N, K, 1.0, dA.data_ptr(), K, dB.data_ptr(), N, 0.0, dC.data_ptr(), N)
    sync()
    assert np.array_equal(host(dC, np.uint16), want)
    want32 = oracle.linear_bf16(A16, np.ascontiguousarray(B16.T), c_f32=True, threads=8)
    dC32 = torch.zeros((M, N), dtype=torch.float32, device="cuda")
    ctx().gemm_bf16(M, N, K, 1.0, dA.data_ptr(), K, dB.data_ptr(), N, 0.0, dC32.data_ptr(), N, capi.FLOAT32)
    sync()
    assert np.array_equal(host(dC32), want32)


def test_hgemm_n_and_k_not_multiples_of_eight(cta_group):
    # only the row pitches have to be 16-byte multiples; N and K themselves may be ragged
    M, N, K, ld = 70, 20, 20, 24
    Ab = oracle.f32_to_f16(dg.int_fill((M, ld), -10, 9, tag=83))
    Bb = oracle.f32_to_f16(dg.int_fill((K, ld), -10, 9, tag=84))
    Cb = oracle.f32_to_f16(dg.int_fill((M, ld), -3, 3, tag=85))
    dA, dB, dC = dev(Ab), dev(Bb), dev(Cb)
    ctx().hgemm_f16(M, N, K, 1.0, dA.data_ptr(), ld, dB.data_ptr(), ld, 1.0, dC.data_ptr(), ld, capi.FLOAT16)
    sync()
    got = host(dC, np.uint16)
    want = oracle.hgemm_f16(np.ascontiguousarray(Ab[:, :K]), np.ascontiguousarray(Bb[:, :N]),
                            np.ascontiguousarray(Cb[:, :N]), 1.0, 1.0)
    assert np.array_equal(got[:, :N], want) and np.array_equal(got[:, N:], Cb[:, N:])


def test_two_scans_on_two_streams_do_not_deadlock():
    # the scan's CTAs wait on each other; the cooperative launch makes concurrent scans take turns
    c1, c2 = ctx(), capi.Context(0)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    x = dg.int_fill(1 << 24, -2, 2, np.int32, tag=86)
    d = dev(x)
    o1, o2 = torch.empty_like(d), torch.empty_like(d)
    torch.cuda.synchronize()
    for _ in range(4):
        c1.scan_i32(x.size, d.data_ptr(), o1.data_ptr(), 1, s1.cuda_stream)
        c2.scan_i32(x.size, d.data_ptr(), o2.data_ptr(), 2, s2.cuda_stream)
    torch.cuda.synchronize()
    assert np.array_equal(host(o1), oracle.scan_i32(x, 1)) and np.array_equal(host(o2), oracle.scan_i32(x, 2))
    c2.close()


@pytest.mark.parametrize("shape", [(64, 64, 4096), (512, 576, 480), (128, 128, 8192), (100, 70, 1033), (256, 256, 256), (1280, 2048, 4096)])
def test_sgemm_split_k_shapes(shape):
    # few tiles, long K: the kernel splits K across CTAs and the last CTA of a tile adds the splits in a
    # fixed order (cuda/gemm_fp32.cu:162-258 is the reference's atomicAdd split-K study; 64x4096x64 is its case)
    M, N, K = shape
    A, B, C = dg.int_fill((M, K), -2, 5, tag=87), dg.int_fill((K, N), -2, 5, tag=88), dg.int_fill((M, N), -2, 5, tag=89)
    for ab in ((1.0, 3.0), (1.0, 0.0)):
        want = oracle.gemm_f32(A, B, C, *ab, order="ikj", threads=8)
        for _ in range(2):  # twice: tickets must be back at rest
            assert np.array_equal(bits32(run_sgemm(A, B, C, *ab)), bits32(want))
    Af, Bf = dg.uniform_fill((M, K), tag=90), dg.uniform_fill((K, N), tag=91)
    got = run_sgemm(Af, Bf, C, 1.0, 0.0)
    want = oracle.gemm_f32(Af, Bf, C, 1.0, 0.0, order="ikj", threads=8)
    assert np.max(np.abs(got - want)) <= 2 * K * EPS
    assert np.array_equal(got, run_sgemm(Af, Bf, C, 1.0, 0.0))  # deterministic: same bits every run


# ------------------------------------------------------------------ wide (256 x 512) tiles
@pytest.mark.parametrize("shape", [(256, 512, 64), (300, 520, 200), (512, 1544, 1024), (1024, 2304, 2000), (1, 8, 8)])
def test_wide_tiles_bit_exact(shape):
    """The tensor-core kernel can cut C into 256 x 512 tiles (one TMEM accumulator, A loaded once per 512 columns);
    the library picks them by a cost model for long problems.  Forced here on small and ragged shapes: fp16 gemm
    (B[K,N]), bf16 / int8 linear (Bt[N,K]) must give the oracle's bits exactly as the 256 x 256 tiles do."""
    M, N, K = shape
    c = ctx()
    c.debug_set("wide_tiles", 2)
    try:
        A16 = oracle.f32_to_f16(dg.int_fill((M, K), -10, 9, tag=61))
        B16 = oracle.f32_to_f16(dg.int_fill((K, N), -10, 9, tag=62))
        C0 = oracle.f32_to_f16(dg.int_fill((M, N), -10, 9, tag=63))
        got = run_hgemm(A16, B16, C0, 2.0, -1.0)
        assert np.array_equal(got, oracle.hgemm_f16(A16, B16, C0, 2.0, -1.0, threads=8))
        Ab = oracle.f32_to_bf16(dg.int_fill((M, K), -8, 8, tag=64))
        Bb = oracle.f32_to_bf16(dg.int_fill((N, K), -8, 8, tag=65))
        dA, dB = dev(Ab), dev(Bb)
        dC = torch.zeros((M, N), dtype=torch.int16, device="cuda")
        c.linear_bf16(M, N, K, dA.data_ptr(), dB.data_ptr(), dC.data_ptr(), capi.BFLOAT16)
        sync()
        assert np.array_equal(host(dC, np.uint16), oracle.linear_bf16(Ab, Bb, threads=8))
        if N % 4 == 0:
            K8 = (K + 15) // 16 * 16
            A8 = dg.int_fill((M, K8), -128, 127, np.int8, tag=66)
            B8 = dg.int_fill((N, K8), -128, 127, np.int8, tag=67)
            dA, dB = dev(A8), dev(B8)
            dC = torch.zeros((M, N), dtype=torch.int32, device="cuda")
            c.linear_s8(M, N, K8, dA.data_ptr(), dB.data_ptr(), dC.data_ptr())
            sync()
            assert np.array_equal(host(dC), oracle.linear_s8(A8, B8, threads=8))
    finally:
        c.debug_set("wide_tiles", 0)
