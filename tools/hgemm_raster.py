"""fp16 GEMM 16384^3: rasterisation group height (tile rows per group) vs sustained rate and clocks."""
import os, sys, ctypes, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hpc_b200 import capi
ctx = capi.Context(0)
L = capi.lib()
n = int(os.environ.get("N", "16384"))
if os.environ.get("KIND", "f16") == "s8":     # int8 linear (KIND=s8 N=8192)
    A = torch.randint(0, 17, (n, n), device="cuda", dtype=torch.int32).to(torch.int8)
    B = torch.randint(-8, 9, (n, n), device="cuda", dtype=torch.int32).to(torch.int8)
    C = torch.zeros(n, n, device="cuda", dtype=torch.int32)
    f = lambda: ctx.linear_s8(n, n, n, A.data_ptr(), B.data_ptr(), C.data_ptr(), torch.cuda.current_stream().cuda_stream)
else:
    A = torch.randint(-10, 10, (n, n), device="cuda").half()
    B = torch.randint(-10, 10, (n, n), device="cuda").half()
    C = torch.zeros(n, n, device="cuda", dtype=torch.float16)
    f = lambda: ctx.hgemm_f16(n, n, n, 1.0, A.data_ptr(), n, B.data_ptr(), n, 0.0, C.data_ptr(), n, capi.FLOAT16, torch.cuda.current_stream().cuda_stream)
for g in [int(x) for x in (sys.argv[1:] or ["8", "4", "6", "12", "16", "8"])]:
    sync = 1 if g < 0 else 0      # a negative group height = that height with the soft wave rendezvous on
    g = abs(g)
    ctx.debug_set("wave_sync", sync)
    ctx.debug_set("group_m", g)
    for _ in range(int(os.environ.get('WARM', '3'))): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = int(os.environ.get('REPS', '60'))
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    clk = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader"], capture_output=True, text=True).stdout.strip()
    print("sync %d group_m %2d  %.3f ms  %.0f TFLOP/s   (after: %s)" % (sync, g, ms, 2.0 * n ** 3 / ms * 1e-9, clk))
