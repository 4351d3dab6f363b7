"""Oracle vs the committed golden vectors (tests/golden/ref_vectors.npz), which were
produced by the reference's own host code (tests/golden/make_golden.py).  Runs
anywhere -- no /root/reference needed."""
import os

import numpy as np

import oracle

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_vectors.npz"))


def test_gemm_golden():
    for name in ("gemm_int", "gemm_flt"):
        got = oracle.gemm_f32(G[name + "_A"], G[name + "_B"], G[name + "_C0"], 1.0, 3.0)
        assert np.array_equal(got.view(np.uint32), G[name + "_C"].view(np.uint32))


def test_hgemm_golden():
    for name in ("hgemm_int", "hgemm_nrm"):
        got = oracle.hgemm_f16(G[name + "_A16"], G[name + "_B16"])
        assert np.array_equal(got, G[name + "_C16"])


def test_linear_golden():
    assert np.array_equal(oracle.linear_s8(G["lin8_A"], G["lin8_Bt"]), G["lin8_C"])
    assert np.array_equal(oracle.linear_bf16(G["linbf_A"], G["linbf_Bt"]), G["linbf_C"])


def test_reduction_golden():
    for name in ("red_int", "red_flt"):
        a, b, w = G[name + "_a"], G[name + "_b"], G[name + "_dot_sum_nrm2"]
        got = np.array([oracle.dot_f32(a, b), oracle.sum_f32(a), oracle.nrm2_f32(a)], np.float32)
        assert np.array_equal(got.view(np.uint32), w.view(np.uint32))


def test_scan_golden():
    x = G["scan_in"]
    assert np.array_equal(oracle.scan_f32(x), G["scan_out_f32"])
    assert np.array_equal(oracle.scan_i32(x.astype(np.int32)).astype(np.float32), G["scan_out_f32"])
