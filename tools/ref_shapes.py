"""The reference's OWN benchmark shapes on the B200 kernels, next to what the reference prints / publishes
for them (other hardware).  Device-resident inputs, CUDA events around 50 launches.

  cux demo (0-frameworks/cux/example/demo.cpp:18,51,77): dot / nrm2 over 4 096 000 floats, gemm 512x576x480
  cuda/gemm_fp32.cu:637-638 "normal test" 1280x2048x4096 and :714-718 split-K case 64x64x4096
  cuda/gemm_fp16_wmma.cu:806-807 2048x2048x4096 (fp16 in, fp16 out)
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from hpc_b200 import capi

ctx = capi.Context(0)
st = lambda: torch.cuda.current_stream().cuda_stream


def timed(f, reps=50):
    for _ in range(5):
        f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        f()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


rows = []
n = 4096000
a = torch.randint(-2, 3, (n,), device="cuda").float()
b = torch.randint(-2, 3, (n,), device="cuda").float()
r = torch.zeros(1, device="cuda")
ms = timed(lambda: ctx.dot_f32(n, a.data_ptr(), b.data_ptr(), r.data_ptr(), st()))
rows.append(("cux dot, 4 096 000 floats", ms, "%.0f GB/s" % (8 * n / ms * 1e-6), "none published"))
ms = timed(lambda: ctx.nrm2_f32(n, a.data_ptr(), r.data_ptr(), st()))
rows.append(("cux nrm2, 4 096 000 floats", ms, "%.0f GB/s" % (4 * n / ms * 1e-6), "none published"))

for (M, N, K), pub in [((512, 576, 480), "cuBLAS sgemm 0.189 ms, best own kernel 0.95 ms (unnamed GPU, demo screenshot)"),
                       ((1280, 2048, 4096), "best kernel v5 8.40 ms (Tesla T4, gemm_fp32.cu:704-712)"),
                       ((64, 64, 4096), "split-K v3 0.352 ms (Tesla T4, gemm_fp32.cu:714-718)")]:
    A = torch.randint(-2, 6, (M, K), device="cuda").float()
    B = torch.randint(-2, 6, (K, N), device="cuda").float()
    C = torch.zeros(M, N, device="cuda")
    ms = timed(lambda: ctx.sgemm_f32(M, N, K, 1.0, A.data_ptr(), K, B.data_ptr(), N, 0.0, C.data_ptr(), N, st()))
    rows.append(("fp32 gemm %dx%dx%d" % (M, N, K), ms, "%.1f TFLOP/s" % (2.0 * M * N * K / ms * 1e-9), pub))

for (M, N, K), pub in [((512, 576, 480), "cublasSgemmEx half 0.155 ms (unnamed GPU, demo screenshot)"),
                       ((2048, 2048, 4096), "wmma v5 0.908 ms (RTX 4050 Laptop), v2 0.460 ms (RTX 3080), gemm_fp16_wmma.cu:871-882")]:
    A = torch.randint(-10, 10, (M, K), device="cuda").half()
    B = torch.randint(-10, 10, (K, N), device="cuda").half()
    C = torch.zeros(M, N, device="cuda", dtype=torch.float16)
    ms = timed(lambda: ctx.hgemm_f16(M, N, K, 1.0, A.data_ptr(), K, B.data_ptr(), N, 0.0, C.data_ptr(), N, capi.FLOAT16, st()))
    rows.append(("fp16 gemm %dx%dx%d (fp32 accumulate)" % (M, N, K), ms, "%.1f TFLOP/s" % (2.0 * M * N * K / ms * 1e-9), pub))

print("%-44s %10s  %-14s  %s" % ("shape", "ms", "rate", "reference's own number (other hardware)"))
for name, ms, rate, pub in rows:
    print("%-44s %10.4f  %-14s  %s" % (name, ms, rate, pub))
print(torch.cuda.get_device_name(0))
