"""Timing of the small / skinny SGEMM shapes the reference studies (split-K case, cux demo shape)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hpc_b200 import capi
ctx = capi.Context(0)
for (M, N, K) in [(64, 64, 4096), (512, 576, 480), (1280, 2048, 4096), (8192, 8192, 8192)]:
    A = torch.randn(M, K, device="cuda"); B = torch.randn(K, N, device="cuda"); C = torch.zeros(M, N, device="cuda")
    f = lambda: ctx.sgemm_f32(M, N, K, 1.0, A.data_ptr(), K, B.data_ptr(), N, 0.0, C.data_ptr(), N, torch.cuda.current_stream().cuda_stream)
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print("sgemm %5dx%5dx%5d  %.4f ms  %.2f TFLOP/s" % (M, N, K, ms, 2.0 * M * N * K / ms * 1e-9))
