"""Small shapes of every kernel, for compute-sanitizer (memcheck / racecheck) runs:
    python tools/sanitize_smoke.py && compute-sanitizer --tool memcheck python tools/sanitize_smoke.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import oracle
from hpc_b200 import capi, datagen as dg

ctx = capi.Context(0)
dev = lambda x: torch.from_numpy(np.ascontiguousarray(x).view(np.int16) if x.dtype == np.uint16 else np.ascontiguousarray(x)).cuda()
n = 100003
a, b = dg.int_fill(n, -2, 2, tag=1), dg.int_fill(n, -2, 2, tag=2)
da, db = dev(a), dev(b)
r = torch.zeros(1, device="cuda")
ctx.dot_f32(n, da.data_ptr(), db.data_ptr(), r.data_ptr()); torch.cuda.synchronize()
assert float(r[0]) == float(oracle.dot_f32(a, b))
ctx.nrm2_f32(n, da.data_ptr(), r.data_ptr()); torch.cuda.synchronize()
x = dg.int_fill(3 * 8192 + 77, -2, 2, np.int32, tag=3)
dx = dev(x); out = torch.empty_like(dx)
ctx.scan_i32(x.size, dx.data_ptr(), out.data_ptr(), 5); torch.cuda.synchronize()
assert np.array_equal(out.cpu().numpy(), oracle.scan_i32(x, 5))
M, N, K = 200, 136, 72
A, B, C = dg.int_fill((M, K), -2, 5, tag=4), dg.int_fill((K, N), -2, 5, tag=5), dg.int_fill((M, N), -2, 5, tag=6)
dA, dB, dC = dev(A), dev(B), dev(C)
ctx.sgemm_f32(M, N, K, 1.0, dA.data_ptr(), K, dB.data_ptr(), N, 3.0, dC.data_ptr(), N); torch.cuda.synchronize()
assert np.array_equal(dC.cpu().numpy(), oracle.gemm_f32(A, B, C, 1.0, 3.0))
A2, B2, C2 = dg.int_fill((256, 128), -2, 5, tag=7), dg.int_fill((128, 256), -2, 5, tag=8), np.zeros((256, 256), np.float32)
dA, dB, dC = dev(A2), dev(B2), dev(C2)
ctx.sgemm_f32(256, 256, 128, 1.0, dA.data_ptr(), 128, dB.data_ptr(), 256, 0.0, dC.data_ptr(), 256); torch.cuda.synchronize()
assert np.array_equal(dC.cpu().numpy(), oracle.gemm_f32(A2, B2, C2, 1.0, 0.0))
A16, B16 = oracle.f32_to_f16(A), oracle.f32_to_f16(B)
dA, dB = dev(A16), dev(B16); dC = torch.zeros((M, N), dtype=torch.int16, device="cuda")
for cg in (2, 1):
    ctx.debug_set("cta_group", cg)
    ctx.hgemm_f16(M, N, K, 1.0, dA.data_ptr(), K, dB.data_ptr(), N, 0.0, dC.data_ptr(), N, capi.FLOAT16); torch.cuda.synchronize()
    assert np.array_equal(dC.cpu().numpy().view(np.uint16), oracle.hgemm_f16(A16, B16))
ctx.debug_set("cta_group", 0)
A8, B8 = dg.int_fill((130, 64), -128, 127, np.int8, tag=9), dg.int_fill((264, 64), -128, 127, np.int8, tag=10)
dA, dB = dev(A8), dev(B8); dC = torch.zeros((130, 264), dtype=torch.int32, device="cuda")
ctx.linear_s8(130, 264, 64, dA.data_ptr(), dB.data_ptr(), dC.data_ptr()); torch.cuda.synchronize()
assert np.array_equal(dC.cpu().numpy(), oracle.linear_s8(A8, B8))
print("sanitize_smoke ok")
