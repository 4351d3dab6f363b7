"""256x256 vs 256x512 tiles of the tensor GEMM: correctness on ragged shapes, then burst / sustained at BASELINE sizes."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hpc_b200 import capi
ctx = capi.Context(0)
st = lambda: torch.cuda.current_stream().cuda_stream
ok = True
for (M, N, K) in [(256, 512, 64), (512, 1024, 256), (300, 520, 200), (1024, 2304, 2000), (1, 8, 8), (2048, 4096, 4096)]:
    A = torch.randint(-10, 10, (M, K), device="cuda").half(); B = torch.randint(-10, 10, (K, N), device="cuda").half()
    Bt = B.t().contiguous()
    A8 = torch.randint(-128, 128, (M, (K + 15) // 16 * 16), device="cuda", dtype=torch.int32).to(torch.int8)
    Bt8 = torch.randint(-128, 128, (N, A8.shape[1]), device="cuda", dtype=torch.int32).to(torch.int8)
    outs = []
    for w in (1, 2):
        ctx.debug_set("wide_tiles", w)
        C = torch.zeros((M, N), dtype=torch.float16, device="cuda"); C2 = torch.zeros_like(C)
        C8 = torch.zeros((M, N), dtype=torch.int32, device="cuda")
        ctx.hgemm_f16(M, N, K, 1.0, A.data_ptr(), K, B.data_ptr(), N, 0.0, C.data_ptr(), N, capi.FLOAT16, st())
        ctx.linear_f16(M, N, K, A.data_ptr(), Bt.data_ptr(), C2.data_ptr(), capi.FLOAT16, st())
        if N % 4 == 0:
            ctx.linear_s8(M, N, A8.shape[1], A8.data_ptr(), Bt8.data_ptr(), C8.data_ptr(), st())
        torch.cuda.synchronize()
        outs.append((C, C2, C8))
    same = all(torch.equal(a, b) for a, b in zip(outs[0], outs[1])) and torch.equal(outs[0][0], outs[0][1])
    print((M, N, K), "wide == narrow:", same, flush=True)
    ok &= same
if not ok:
    sys.exit(1)
def bench(name, f, flop):
    for w in (1, 2, 1, 2):
        ctx.debug_set("wide_tiles", w)
        for _ in range(3): f()
        torch.cuda.synchronize(); time.sleep(1.0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): f()
        e1.record(); torch.cuda.synchronize()
        burst = e0.elapsed_time(e1) / 10
        reps = max(20, int(2000 / burst))
        e0.record()
        for _ in range(reps): f()
        e1.record(); torch.cuda.synchronize()
        sus = e0.elapsed_time(e1) / reps
        print("%-22s tiles 256x%d  burst %.3f ms %.0f T  sustained %.3f ms %.0f T" % (name, 256 * w, burst, flop / burst * 1e-9, sus, flop / sus * 1e-9), flush=True)
n = 16384
for data in ("int", "normal"):
    A = (torch.randint(-10, 10, (n, n), device="cuda") if data == "int" else torch.randn(n, n, device="cuda")).half()
    B = (torch.randint(-10, 10, (n, n), device="cuda") if data == "int" else torch.randn(n, n, device="cuda")).half()
    C = torch.zeros(n, n, device="cuda", dtype=torch.float16)
    bench("hgemm 16384^3 " + data, lambda: ctx.hgemm_f16(n, n, n, 1.0, A.data_ptr(), n, B.data_ptr(), n, 0.0, C.data_ptr(), n, capi.FLOAT16, st()), 2.0 * n ** 3)
del A, B, C
n = 8192
A8 = torch.randint(0, 17, (n, n), device="cuda", dtype=torch.int32).to(torch.int8); B8 = torch.randint(-8, 9, (n, n), device="cuda", dtype=torch.int32).to(torch.int8)
C8 = torch.zeros(n, n, device="cuda", dtype=torch.int32)
bench("linear_s8 8192^3", lambda: ctx.linear_s8(n, n, n, A8.data_ptr(), B8.data_ptr(), C8.data_ptr(), st()), 2.0 * n ** 3)
