"""Time the copying (self-pull) and the plain fp16 GEMM with the library named by CUXB_LIB_PATH (one GPU)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hpc_b200 import capi
import ctypes
_L = ctypes.CDLL(capi.LIB_PATH)
capi.SIGNATURES = {k: v for k, v in capi.SIGNATURES.items() if hasattr(_L, k)}   # older builds lack the newest symbols
ctx = capi.Context(0)
M, N, K = 8192, 16384, 16384
A = torch.randint(-10, 10, (M, K), device="cuda").half(); B = torch.randint(-10, 10, (K, N), device="cuda").half()
Bl = torch.zeros_like(B); C = torch.zeros((M, N), dtype=torch.float16, device="cuda")
st = torch.cuda.current_stream().cuda_stream
def t(f, reps=10):
    best = 1e9
    for _ in range(3):
        for _ in range(2): f()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps): f()
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / reps)
    return best
try: ctx.debug_set("wide_tiles", 1)
except Exception: pass
plain = t(lambda: ctx.hgemm_f16(M, N, K, 1.0, A.data_ptr(), K, B.data_ptr(), N, 0.0, C.data_ptr(), N, capi.FLOAT16, st))
pull = t(lambda: ctx.mg_hgemm_f16(M, N, K, 1.0, A.data_ptr(), K, B.data_ptr(), Bl.data_ptr(), N, 0.0, C.data_ptr(), N, False, capi.FLOAT16, st))
print(os.path.basename(os.environ.get("CUXB_LIB_PATH", "HEAD")), "plain narrow %.3f ms  self-pull %.3f ms  B ok %s" % (plain, pull, bool(torch.equal(Bl, B))), flush=True)
