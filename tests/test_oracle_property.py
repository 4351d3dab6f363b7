"""Property tests (hypothesis) pinning oracle/oracle.c to the reference's own host code on RANDOM shapes and
values -- ragged sizes, odd leading cases, negative alpha / beta, values that round -- beyond the fixed cases
of test_oracle_vs_ref.py.  Bit-for-bit, as there."""
import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

import oracle

ref = oracle.ref()
pytestmark = pytest.mark.skipif(ref is None, reason="oracle/_ref not built (no /root/reference)")
CFG = dict(max_examples=40, deadline=None)


def _f32(rng, shape):
    # a mix of magnitudes so that sums round differently depending on their order
    return (rng.standard_normal(shape) * 10.0 ** rng.integers(-3, 4, size=shape)).astype(np.float32)


def bits(x):
    return np.ascontiguousarray(x).view(np.uint32 if x.dtype == np.float32 else x.dtype)


@settings(**CFG)
@given(st.integers(1, 40), st.integers(1, 40), st.integers(1, 70), st.sampled_from([1.0, -1.0, 0.5, 3.0]),
       st.sampled_from([0.0, 1.0, -2.0, 3.0]), st.integers(0, 2 ** 31 - 1))
def test_gemm_v0_random_shapes(M, N, K, alpha, beta, seed):
    rng = np.random.default_rng(seed)
    A, B, C = _f32(rng, (M, K)), _f32(rng, (K, N)), _f32(rng, (M, N))
    want = C.copy()
    ref.ref_cux_gemm(0, M, N, K, alpha, A, K, B, N, beta, want, N)
    assert np.array_equal(bits(oracle.gemm_f32(A, B, C, alpha, beta, order="v0")), bits(want))


@settings(**CFG)
@given(st.integers(1, 5000), st.integers(0, 2 ** 31 - 1))
def test_dot_sum_nrm2_random_lengths(n, seed):
    rng = np.random.default_rng(seed)
    a, b = _f32(rng, n), _f32(rng, n)
    out = np.zeros(1, np.float32)
    ref.ref_cux_dot_v0(n, a, b, out)
    assert bits(out)[0] == bits(np.float32(oracle.dot_f32(a, b)).reshape(1))[0]
    ref.ref_cux_nrm2_v0(n, a, out)
    assert bits(out)[0] == bits(np.float32(oracle.nrm2_f32(a)).reshape(1))[0]
    ref.ref_reduce_host(a, out, n)
    assert bits(out)[0] == bits(np.float32(oracle.sum_f32(a)).reshape(1))[0]


@settings(**CFG)
@given(st.integers(1, 24), st.integers(1, 24), st.integers(1, 200), st.integers(0, 2 ** 31 - 1))
def test_linear_int8_random_shapes(M, N, K, seed):
    rng = np.random.default_rng(seed)
    A = rng.integers(-128, 128, size=(M, K)).astype(np.int8)
    Bt = rng.integers(-128, 128, size=(N, K)).astype(np.int8)
    want = np.zeros((M, N), np.int32)
    ref.ref_linear_s8(0, M, N, K, A, Bt, want)          # LinearNormalInt8
    assert np.array_equal(oracle.linear_s8(A, Bt), want)


@settings(**CFG)
@given(st.integers(1, 3000), st.integers(0, 2 ** 31 - 1))
def test_scan_f32_random_lengths(n, seed):
    rng = np.random.default_rng(seed)
    x = _f32(rng, n)
    want = np.empty(n, np.float32)
    ref.ref_scan_f32(0, x, n, want)                      # VectorScanNormal
    assert np.array_equal(bits(oracle.scan_f32(x)), bits(want))


@settings(**CFG)
@given(st.integers(1, 24), st.integers(1, 24), st.integers(1, 120), st.integers(0, 2 ** 31 - 1))
def test_hgemm_and_bf16_linear_random_shapes(M, N, K, seed):
    rng = np.random.default_rng(seed)
    A, B = _f32(rng, (M, K)), _f32(rng, (K, N))
    A16, B16 = oracle.f32_to_f16(A), oracle.f32_to_f16(B)    # includes overflow to inf and subnormals
    want = np.zeros((M, N), np.uint16)
    ref.ref_hgemm_host(M, N, K, A16, K, B16, N, want, N)     # GemmHostV1 (half), gemm_fp16_wmma.cu:75-90
    got = oracle.hgemm_f16(A16, B16)
    nan_w, nan_g = (want & 0x7fff) > 0x7c00, (got & 0x7fff) > 0x7c00   # inf - inf: any NaN payload is a NaN
    assert np.array_equal(nan_w, nan_g) and np.array_equal(got[~nan_g], want[~nan_w])
    Ab, Bb = oracle.f32_to_bf16(A), oracle.f32_to_bf16(np.ascontiguousarray(B.T))
    wb = np.zeros((M, N), np.uint16)
    assert ref.ref_linear_bf16(0, M, N, K, Ab, Bb, wb) == 0  # LinearNormalBf16, linear.cpp:137-151
    gb = oracle.linear_bf16(Ab, Bb)
    nan_w, nan_g = (wb & 0x7fff) > 0x7f80, (gb & 0x7fff) > 0x7f80
    assert np.array_equal(nan_w, nan_g) and np.array_equal(gb[~nan_g], wb[~nan_w])
