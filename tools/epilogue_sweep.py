"""Tensor GEMM: direct-store vs TMA-store epilogue x 256x256 vs 256x512 tiles.  Median of 3 alternations."""
import os, sys
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
for (M, N, K) in [(4096, 4096, 4096), (8192, 8192, 8192), (16384, 16384, 2048), (8192, 8192, 1024), (8192, 8192, 2048), (16384, 16384, 16384),
                  (2048, 16384, 16384), (6144, 6144, 6144), (512, 576, 480), (2048, 2048, 4096)]:
    N8 = N // 8 * 8; K8 = K // 16 * 16
    A = torch.randint(-10, 10, (M, K8), device="cuda").half(); B = torch.randint(-10, 10, (K8, N8), device="cuda").half()
    C = torch.zeros(M, N8, device="cuda", dtype=torch.float16)
    A8 = torch.randint(-8, 9, (M, K8), device="cuda", dtype=torch.int32).to(torch.int8); B8 = torch.randint(-8, 9, (N8, K8), device="cuda", dtype=torch.int32).to(torch.int8)
    C8 = torch.zeros(M, N8, device="cuda", dtype=torch.int32)
    cfgs = [(1, 1), (1, 0), (2, 1), (2, 0)]   # (tile width, direct epilogue)
    r = {(k, c): [] for k in ("f16", "s8") for c in cfgs}
    ref = {}
    for rep in range(3):
        for c in cfgs:
            ctx.debug_set("wide_tiles", c[0]); ctx.debug_set("direct_epilogue", 1 if c[1] else 2)   # 1 = direct stores, 2 = staged wherever possible
            r[("f16", c)].append(t(lambda: ctx.hgemm_f16(M, N8, K8, 1.0, A.data_ptr(), K8, B.data_ptr(), N8, 0.0, C.data_ptr(), N8, capi.FLOAT16, st())))
            r[("s8", c)].append(t(lambda: ctx.linear_s8(M, N8, K8, A8.data_ptr(), B8.data_ptr(), C8.data_ptr(), st())))
            if rep == 0:
                torch.cuda.synchronize()
                if not ref: ref = {"f16": C.clone(), "s8": C8.clone()}
                elif not (torch.equal(C, ref["f16"]) and torch.equal(C8, ref["s8"])):
                    print("MISMATCH", (M, N, K), c, flush=True); sys.exit(1)
    m = {k: sorted(v)[1] for k, v in r.items()}
    fl = 2.0 * M * N8 * K8
    line = "%-22s" % ((M, N, K),)
    for k in ("f16", "s8"):
        line += "  %s:" % k + " ".join(" w%d%s %.4f (%4.0fT)" % (c[0], "d" if c[1] else "t", m[(k, c)], fl / m[(k, c)] * 1e-9) for c in cfgs)
    print(line, flush=True)
    del A, B, C, A8, B8, C8
