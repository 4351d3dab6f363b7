"""Where do 256x512 tiles pay?  fp16 GEMM and int8 linear over a range of shapes, 256x256 vs 256x512, best of 3 x 20 launches."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hpc_b200 import capi
ctx = capi.Context(0)
st = lambda: torch.cuda.current_stream().cuda_stream
def t(f, reps=20):
    best = 1e9
    for _ in range(2):
        for _ in range(2): f()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps): f()
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / reps)
    return best
for (M, N, K) in [(4096, 4096, 4096), (8192, 8192, 8192), (16384, 16384, 2048), (16384, 16384, 1024), (8192, 8192, 1024), (8192, 8192, 2048),
                  (8192, 8192, 4096), (16384, 16384, 16384), (1024, 16384, 16384), (2048, 16384, 16384), (16384, 2048, 16384), (4096, 4096, 16384),
                  (8192, 16384, 8192), (6144, 6144, 6144)]:
    N8 = N // 8 * 8; K8 = K // 16 * 16
    A = torch.randint(-10, 10, (M, K8), device="cuda").half(); B = torch.randint(-10, 10, (K8, N8), device="cuda").half()
    C = torch.zeros(M, N8, device="cuda", dtype=torch.float16)
    A8 = torch.randint(-8, 9, (M, K8), device="cuda", dtype=torch.int32).to(torch.int8); B8 = torch.randint(-8, 9, (N8, K8), device="cuda", dtype=torch.int32).to(torch.int8)
    C8 = torch.zeros(M, N8, device="cuda", dtype=torch.int32)
    r = {"f16_1": [], "f16_2": [], "s8_1": [], "s8_2": []}
    for rep in range(3):          # alternate, keep the median: the power state drifts
        for w in (1, 2):
            ctx.debug_set("wide_tiles", w)
            r["f16_%d" % w].append(t(lambda: ctx.hgemm_f16(M, N8, K8, 1.0, A.data_ptr(), K8, B.data_ptr(), N8, 0.0, C.data_ptr(), N8, capi.FLOAT16, st())))
            r["s8_%d" % w].append(t(lambda: ctx.linear_s8(M, N8, K8, A8.data_ptr(), B8.data_ptr(), C8.data_ptr(), st())))
    r = {k: sorted(v)[1] for k, v in r.items()}
    fl = 2.0 * M * N8 * K8
    print("%-22s fp16 %.4f -> %.4f ms (%+.1f%%, %.0f T)   int8 %.4f -> %.4f ms (%+.1f%%, %.0f T)" % (
        (M, N, K), r["f16_1"], r["f16_2"], (r["f16_1"] / r["f16_2"] - 1) * 100, fl / r["f16_2"] * 1e-9,
        r["s8_1"], r["s8_2"], (r["s8_1"] / r["s8_2"] - 1) * 100, fl / r["s8_2"] * 1e-9), flush=True)
    del A, B, C, A8, B8, C8
