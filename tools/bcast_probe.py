"""Probe: how fast does the in-kernel B pull run when there is (almost) no compute to hide behind?
Single process, 2 GPUs, peer access.  python tools/bcast_probe.py"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hpc_b200 import capi
L = capi.lib()
for a in (0, 1):
    capi.check(L.cuxb_enable_peer_access(a, 1 - a))
c0, c1 = capi.Context(0), capi.Context(1)
N = K = 16384
for M in (256, 2048, 16384):
    B0 = torch.ones((K, N), dtype=torch.float16, device="cuda:0")
    B1 = torch.zeros((K, N), dtype=torch.float16, device="cuda:1")
    A1 = torch.ones((M, K), dtype=torch.float16, device="cuda:1")
    C1 = torch.zeros((M, N), dtype=torch.float16, device="cuda:1")
    torch.cuda.set_device(1)
    c1.hgemm_f16(M, N, K, 1.0, A1.data_ptr(), K, B1.data_ptr(), N, 0.0, C1.data_ptr(), N)  # load module
    torch.cuda.synchronize(1)
    for mode in ("resident", "pull"):
        ts = []
        for rep in range(6):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            if mode == "pull":
                c1.mg_hgemm_f16(M, N, K, 1.0, A1.data_ptr(), K, B0.data_ptr(), B1.data_ptr(), N, 0.0, C1.data_ptr(), N, False)
            else:
                c1.hgemm_f16(M, N, K, 1.0, A1.data_ptr(), K, B1.data_ptr(), N, 0.0, C1.data_ptr(), N)
            e1.record()
            torch.cuda.synchronize(1)
            ts.append(e0.elapsed_time(e1))
        t = sorted(ts)[len(ts) // 2]
        print("M=%5d %-8s %.3f ms  (pull 512 MiB -> %.0f GB/s if pull-bound)" % (M, mode, t, 0.537 / (t * 1e-3)))
