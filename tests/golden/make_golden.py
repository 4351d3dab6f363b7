#!/usr/bin/env python3
"""Generate tests/golden/*.npz from the REFERENCE's own host code.

Run in the authoring container (needs /root/reference -> oracle/_ref/libcuxref.so):

    python oracle/build.py && python tests/golden/make_golden.py

Each fixture stores the seeded inputs and the outputs produced by the reference
functions named in its key, so the oracle (and the GPU path) can be pinned to
them on machines where the reference tree does not exist.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from hpc_b200 import datagen as dg  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def main():
    ref = oracle.ref()
    assert ref is not None, "build oracle/_ref first"
    g = {}
    # --- cux gemm fp32 (GemmHostV0, gemm.cpp:13-34), demo-style alpha=1 beta=3 (demo.cpp:77,97)
    for name, (M, N, K), fill in [("gemm_int", (96, 128, 64), "int"), ("gemm_flt", (64, 48, 80), "flt")]:
        if fill == "int":
            A, B, C = dg.int_fill((M, K), -2, 5, tag=101), dg.int_fill((K, N), -2, 5, tag=102), dg.int_fill((M, N), -2, 5, tag=103)
        else:
            A, B, C = dg.uniform_fill((M, K), tag=101), dg.uniform_fill((K, N), tag=102), dg.uniform_fill((M, N), tag=103)
        out = C.copy()
        ref.ref_cux_gemm(0, M, N, K, 1.0, A, K, B, N, 3.0, out, N)
        g[name + "_A"], g[name + "_B"], g[name + "_C0"], g[name + "_C"] = A, B, C, out
    # --- fp16 gemm fp32-accumulate (GemmHostV1 half, gemm_fp16_wmma.cu:75-90)
    for name, fill in [("hgemm_int", "int"), ("hgemm_nrm", "nrm")]:
        M, N, K = 128, 128, 64
        A = dg.int_fill((M, K), -10, 9, tag=111) if fill == "int" else dg.normal_fill((M, K), tag=111)
        B = dg.int_fill((K, N), -10, 9, tag=112) if fill == "int" else dg.normal_fill((K, N), tag=112)
        A16, B16 = oracle.f32_to_f16(A), oracle.f32_to_f16(B)
        out = np.zeros((M, N), np.uint16)
        ref.ref_hgemm_host(M, N, K, A16, K, B16, N, out, N)
        g[name + "_A16"], g[name + "_B16"], g[name + "_C16"] = A16, B16, out
    # --- int8 linear (LinearNormalInt8, linear.cpp:199-211) and bf16 linear (:137-151)
    M, N, K = 128, 128, 128
    A8 = dg.int_fill((M, K), -128, 127, np.int8, tag=121)
    B8 = dg.int_fill((N, K), -128, 127, np.int8, tag=122)
    C32 = np.empty((M, N), np.int32)
    ref.ref_linear_s8(0, M, N, K, A8, B8, C32)
    g["lin8_A"], g["lin8_Bt"], g["lin8_C"] = A8, B8, C32
    Ab, Bb = oracle.f32_to_bf16(dg.normal_fill((M, K), tag=123)), oracle.f32_to_bf16(dg.normal_fill((N, K), tag=124))
    Cb = np.empty((M, N), np.uint16)
    ref.ref_linear_bf16(0, M, N, K, Ab, Bb, Cb)
    g["linbf_A"], g["linbf_Bt"], g["linbf_C"] = Ab, Bb, Cb
    # --- reductions (DotHostV0 dot_product.cpp:7-13, ReduceHost reduce_fp32.cu:26-31, Nrm2HostV0 nrm2.cpp:7-16)
    n = 40960
    for name, a, b in [("red_int", dg.int_fill(n, -2, 2, tag=131), dg.int_fill(n, -2, 2, tag=132)),
                       ("red_flt", dg.uniform_fill(n, tag=131), dg.uniform_fill(n, tag=132))]:
        w = np.zeros(3, np.float32)
        ref.ref_cux_dot_v0(n, a, b, w[0:1])
        ref.ref_reduce_host(a, w[1:2], n)
        ref.ref_cux_nrm2_v0(n, a, w[2:3])
        g[name + "_a"], g[name + "_b"], g[name + "_dot_sum_nrm2"] = a, b, w
    # --- scan (VectorScanNormal, vector_scan.cpp:21-26), integer-valued so int32 == fp32
    x = dg.int_fill(32768, -2, 2, tag=141)
    s = np.empty_like(x)
    ref.ref_scan_f32(0, x, x.size, s)
    g["scan_in"], g["scan_out_f32"] = x, s
    np.savez_compressed(os.path.join(OUT, "ref_vectors.npz"), **g)
    print("wrote", os.path.join(OUT, "ref_vectors.npz"), sum(v.nbytes for v in g.values()), "bytes raw")


if __name__ == "__main__":
    main()
