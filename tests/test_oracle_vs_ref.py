"""Pin the CPU restatement (oracle/oracle.c) to the reference's OWN host code.

oracle/_ref/libcuxref.so is the reference compiled from /root/reference
(oracle/build.py).  Every oracle function must agree with it BIT FOR BIT on the
same seeded inputs -- integer-valued fills (where any order is exact) AND
non-integer float fills (where only the identical order/rounding agrees).
"""
import numpy as np
import pytest

import oracle
from hpc_b200 import datagen as dg

ref = oracle.ref()
needs_ref = pytest.mark.skipif(ref is None, reason="oracle/_ref not built (no /root/reference)")


def bits(x):
    return np.ascontiguousarray(x).view(np.uint32 if x.dtype == np.float32 else x.dtype)


@needs_ref
@pytest.mark.parametrize("fill", ["int", "float"])
@pytest.mark.parametrize("shape", [(64, 96, 80), (33, 17, 29), (128, 128, 128)])
def test_gemm_v0_bit_exact_vs_cux(fill, shape):
    M, N, K = shape
    if fill == "int":
        A, B, C = dg.int_fill((M, K), -2, 5, tag=1), dg.int_fill((K, N), -2, 5, tag=2), dg.int_fill((M, N), -2, 5, tag=3)
    else:
        A, B, C = dg.uniform_fill((M, K), tag=1), dg.uniform_fill((K, N), tag=2), dg.uniform_fill((M, N), tag=3)
    for alpha, beta in [(1.0, 3.0), (1.0, 0.0), (2.0, -1.0)]:
        want = C.copy()
        ref.ref_cux_gemm(0, M, N, K, alpha, A, K, B, N, beta, want, N)
        got = oracle.gemm_f32(A, B, C, alpha, beta, order="v0")
        assert np.array_equal(bits(got), bits(want))
        # V1 (ikj) has the same per-element order as V0 -> also bit-identical
        want1 = C.copy()
        ref.ref_cux_gemm(1, M, N, K, alpha, A, K, B, N, beta, want1, N)
        got1 = oracle.gemm_f32(A, B, C, alpha, beta, order="ikj", threads=4)
        assert np.array_equal(bits(got1), bits(want1))
        assert np.array_equal(bits(want1), bits(want))


@needs_ref
def test_all_cux_host_versions_agree_on_integer_fill():
    # with integer inputs and partial sums < 2^24 every order is exact.
    M, N, K = 64, 64, 96
    A, B, C = dg.int_fill((M, K), -2, 5, tag=4), dg.int_fill((K, N), -2, 5, tag=5), dg.int_fill((M, N), -2, 5, tag=6)
    want = oracle.gemm_f32(A, B, C, 1.0, 3.0)
    for v in range(6):
        out = C.copy()
        ref.ref_cux_gemm(v, M, N, K, 1.0, A, K, B, N, 3.0, out, N)
        assert np.array_equal(out, want), "GemmHostV%d" % v


@needs_ref
def test_openmp_gemm_matches_ikj_oracle():
    M, N, K = 50, 70, 90
    A, B = dg.uniform_fill((M, K), tag=7), dg.uniform_fill((K, N), tag=8)
    for par in (0, 1):
        out = np.full((M, N), 7.0, np.float32)
        ref.ref_omp_gemm(par, M, N, K, 1.5, A, K, B, N, out, N)
        got = oracle.gemm_f32(A, B, np.zeros((M, N), np.float32), 1.5, 0.0, order="ikj", threads=3)
        assert np.array_equal(bits(out), bits(got))


@needs_ref
@pytest.mark.parametrize("n", [1, 2, 7, 8, 9, 1000, 4096, 100003])
def test_dot_sum_nrm2_bit_exact(n):
    for fill in ("int", "float"):
        a = dg.int_fill(n, -2, 2, tag=n) if fill == "int" else dg.uniform_fill(n, tag=n)
        b = dg.int_fill(n, -2, 2, tag=n + 1) if fill == "int" else dg.uniform_fill(n, tag=n + 1)
        w = np.zeros(1, np.float32)
        ref.ref_cux_dot_v0(n, a, b, w)
        assert bits(np.float32(oracle.dot_f32(a, b))) == bits(w)[0]
        ref.ref_reduce_host(a, w, n)
        assert bits(np.float32(oracle.sum_f32(a))) == bits(w)[0]
        ref.ref_cux_nrm2_v0(n, a, w)
        assert bits(np.float32(oracle.nrm2_f32(a))) == bits(w)[0]


@needs_ref
def test_nrm2_edge_cases():
    w = np.zeros(1, np.float32)
    x = np.array([-3.5], np.float32)
    ref.ref_cux_nrm2_v0(1, x, w)
    assert oracle.nrm2_f32(x) == w[0] == np.float32(3.5)


@needs_ref
@pytest.mark.parametrize("shape", [(16, 24, 32), (40, 30, 50)])
def test_hgemm_bit_exact_vs_wmma_sample_host(shape):
    M, N, K = shape
    for fill in ("int", "normal"):
        if fill == "int":
            A, B = dg.int_fill((M, K), -10, 9, tag=11), dg.int_fill((K, N), -10, 9, tag=12)
        else:
            A, B = dg.normal_fill((M, K), tag=11), dg.normal_fill((K, N), tag=12)
        A16, B16 = oracle.f32_to_f16(A), oracle.f32_to_f16(B)
        want = np.zeros((M, N), np.uint16)
        ref.ref_hgemm_host(M, N, K, A16, K, B16, N, want, N)
        got = oracle.hgemm_f16(A16, B16, threads=2)
        assert np.array_equal(got, want)


def test_f16_converter_matches_numpy_rne():
    # every fp16 bit pattern round-trips; fp32->fp16 equals numpy's RNE on a dense sample
    allh = np.arange(65536, dtype=np.uint16)
    f = oracle.f16_to_f32(allh)
    npf = allh.view(np.float16).astype(np.float32)
    ok = ~np.isnan(npf)
    assert np.array_equal(f[ok], npf[ok])
    assert np.array_equal(oracle.f32_to_f16(f[ok]), allh[ok])
    x = np.concatenate([dg.normal_fill(200000, tag=13) * s for s in (1e-6, 1e-3, 1.0, 300.0, 7e4)])
    with np.errstate(over="ignore"):
        assert np.array_equal(oracle.f32_to_f16(x), x.astype(np.float16).view(np.uint16))


@needs_ref
@pytest.mark.parametrize("shape", [(20, 30, 40), (7, 5, 130)])
def test_linear_int8_and_bf16_bit_exact(shape):
    M, N, K = shape
    A = dg.int_fill((M, K), -128, 127, np.int8, tag=21)
    B = dg.int_fill((N, K), -128, 127, np.int8, tag=22)
    want = np.empty((M, N), np.int32)
    assert ref.ref_linear_s8(0, M, N, K, A, B, want) == 0
    assert np.array_equal(oracle.linear_s8(A, B, threads=2), want)
    # the VNNI variant is u8 x s8: agrees only for A >= 0
    Apos = dg.linear_sample_fill(M, K, np.int8)
    w2 = np.empty((M, N), np.int32)
    if ref.ref_linear_s8(1, M, N, K, Apos, B, w2) == 0:
        assert np.array_equal(oracle.linear_s8(Apos, B), w2)
    Af, Bf = dg.normal_fill((M, K), tag=23), dg.normal_fill((N, K), tag=24)
    A16, B16 = oracle.f32_to_bf16(Af), oracle.f32_to_bf16(Bf)
    wb = np.empty((M, N), np.uint16)
    assert ref.ref_linear_bf16(0, M, N, K, A16, B16, wb) == 0
    assert np.array_equal(oracle.linear_bf16(A16, B16), wb)


@needs_ref
@pytest.mark.parametrize("n", [8, 64, 1000 // 8 * 8, 65536])
def test_scan_matches_vector_scan_sample(n):
    x = dg.int_fill(n, -2, 2, tag=31)
    want = np.empty(n, np.float32)
    ref.ref_scan_f32(0, x, n, want)
    assert np.array_equal(oracle.scan_f32(x), want)
    # int32 restatement == fp32 reference while |prefix| < 2^24 (integers exact)
    xi = x.astype(np.int32)
    assert np.array_equal(oracle.scan_i32(xi).astype(np.float32), want)
    # SIMD variants agree on integer data too
    w2 = np.empty(n, np.float32)
    ref.ref_scan_f32(2, x, n, w2)
    assert np.array_equal(w2, want)


def test_scan_i32_wraparound_matches_numpy():
    x = np.full(1 << 20, 3000, np.int32)  # crosses 2^31
    got = oracle.scan_i32(x, seed=5)
    want = (np.cumsum(x.astype(np.int64)) + 5).astype(np.uint64).astype(np.uint32).view(np.int32)
    assert np.array_equal(got, want)
    assert got.min() < 0  # it really wrapped
    assert oracle.sum_i32(x) == want[-1] - 5


def test_reference_all_cores_dot_matches_the_serial_one_on_integer_fills():
    """std/alg_vector_dot_product.cpp:19-56 (DotProductMTv1, timed by bench.py beside DotHostV0): with the
    cux integer fill every partial sum is exact, so slicing the vector over threads cannot change the result."""
    R = oracle.ref()
    if R is None or not hasattr(R, "ref_std_dot_mt"):
        pytest.skip("oracle/_ref not built with the std/ sample")
    rng = np.random.default_rng(3)
    for n in (1, 7, 1000, 123457):
        a = rng.integers(-2, 3, size=n).astype(np.float32)
        b = rng.integers(-2, 3, size=n).astype(np.float32)
        for th in (1, 3, 8):
            out = np.zeros(1, np.float32)
            R.ref_std_dot_mt(n, a, b, th, out)
            assert out[0] == oracle.dot_f32(a, b), (n, th)
