"""GPU: the single-GPU entry points recorded into a CUDA graph and replayed on fresh data.

A launch that is replayed runs with exactly the arguments it was recorded with, so no protocol state may
live in them: the scan's epoch is counted on the device, the reductions' and split-K tickets reset
themselves.  Every replay must give the oracle's bits for the data that is in the buffers at that moment
(include/cuxb.h lists what may be captured)."""
import numpy as np
import pytest
import torch

import oracle
from hpc_b200 import capi
from hpc_b200 import datagen as dg
from gpu_util import ctx, host

pytestmark = pytest.mark.gpu


def _capture(fn):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(2):          # warm-up outside the capture: workspaces are allocated, modules loaded
            fn(s.cuda_stream)
    s.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        fn(torch.cuda.current_stream().cuda_stream)
    return g


def test_scan_replays_with_a_fresh_epoch_each_time():
    n = 3_000_000 + 77
    x = torch.zeros(n, dtype=torch.int32, device="cuda")
    out = torch.zeros_like(x)
    out2 = torch.zeros_like(x)
    c = ctx()
    L = capi.lib()
    g1 = _capture(lambda st: c.scan_i32(n, x.data_ptr(), out.data_ptr(), 5, st))
    g2 = _capture(lambda st: capi.check(L.cuxb_scan_i32_noalloc(st, n, x.data_ptr(), out2.data_ptr(), -9),
                                        "cuxb_scan_i32_noalloc"))
    for rep in range(4):
        xin = dg.int_fill(n, -1000, 1000, np.int32, tag=200 + rep)
        xin[::3] = 2 ** 30   # wrap-around, and tile aggregates that differ wildly from replay to replay
        x.copy_(torch.from_numpy(xin))
        out.zero_(), out2.zero_()
        g1.replay()
        g2.replay()
        torch.cuda.synchronize()
        assert np.array_equal(host(out), oracle.scan_i32(xin, 5)), rep
        assert np.array_equal(host(out2), oracle.scan_i32(xin, -9)), rep
    # an ordinary call after the replays continues the same epoch sequence
    c.scan_i32(n, x.data_ptr(), out.data_ptr(), 0)
    torch.cuda.synchronize()
    assert np.array_equal(host(out), oracle.scan_i32(xin, 0))


def test_reductions_and_gemms_replay():
    c = ctx()
    n = 1_000_003
    a = torch.zeros(n, device="cuda")
    b = torch.zeros(n, device="cuda")
    res = torch.zeros(3, device="cuda")
    M, N, K = 64, 64, 4096          # split-K shape: per-tile tickets and parked accumulators
    A = torch.zeros((M, K), device="cuda")
    B = torch.zeros((K, N), device="cuda")
    C = torch.zeros((M, N), device="cuda")
    A16 = torch.zeros((256, 512), dtype=torch.float16, device="cuda")
    B16 = torch.zeros((512, 384), dtype=torch.float16, device="cuda")
    C16 = torch.zeros((256, 384), dtype=torch.float16, device="cuda")

    def work(st):
        c.dot_f32(n, a.data_ptr(), b.data_ptr(), res.data_ptr(), st)
        c.sum_f32(n, a.data_ptr(), res.data_ptr() + 4, st)
        c.nrm2_f32(n, b.data_ptr(), res.data_ptr() + 8, st)
        c.sgemm_f32(M, N, K, 1.0, A.data_ptr(), K, B.data_ptr(), N, 0.0, C.data_ptr(), N, st)
        c.hgemm_f16(256, 384, 512, 1.0, A16.data_ptr(), 512, B16.data_ptr(), 384, 0.0, C16.data_ptr(), 384, capi.FLOAT16, st)
    g = _capture(work)
    for rep in range(3):
        ha, hb = dg.int_fill(n, -2, 2, tag=210 + rep), dg.int_fill(n, -2, 2, tag=220 + rep)
        hA, hB = dg.int_fill((M, K), -2, 5, tag=230 + rep), dg.int_fill((K, N), -2, 5, tag=240 + rep)
        hA16 = oracle.f32_to_f16(dg.int_fill((256, 512), -10, 9, tag=250 + rep))
        hB16 = oracle.f32_to_f16(dg.int_fill((512, 384), -10, 9, tag=260 + rep))
        a.copy_(torch.from_numpy(ha)), b.copy_(torch.from_numpy(hb))
        A.copy_(torch.from_numpy(hA)), B.copy_(torch.from_numpy(hB))
        A16.copy_(torch.from_numpy(hA16.view(np.int16)).view(torch.float16))
        B16.copy_(torch.from_numpy(hB16.view(np.int16)).view(torch.float16))
        g.replay()
        torch.cuda.synchronize()
        r = host(res)
        assert r[0] == oracle.dot_f32(ha, hb) and r[1] == oracle.sum_f32(ha) and r[2] == oracle.nrm2_f32(hb), rep
        want = oracle.gemm_f32(hA, hB, np.zeros((M, N), np.float32), 1.0, 0.0, order="ikj", threads=8)
        assert np.array_equal(host(C), want), rep
        assert np.array_equal(host(C16.view(torch.int16), np.uint16), oracle.hgemm_f16(hA16, hB16, threads=8)), rep


def test_graph_recorded_before_a_workspace_grew_still_replays():
    """A recorded scan holds the address of the context's tile-state workspace.  A later, larger scan on the
    same context needs a bigger one; the old one is kept (not freed) until the context is destroyed, so the
    graph recorded before stays valid."""
    c = capi.Context(0)                      # fresh context: its scan workspace starts at 4096 tiles (2^25 elements)
    n = 1_000_003
    x = torch.zeros(n, dtype=torch.int32, device="cuda")
    out = torch.zeros_like(x)
    g = _capture(lambda st: c.scan_i32(n, x.data_ptr(), out.data_ptr(), 3, st))
    big = (1 << 25) + 81920 + 5              # forces the workspace to grow
    gen = torch.Generator(device="cuda")
    gen.manual_seed(3)
    xb = torch.randint(-2, 3, (big,), dtype=torch.int32, device="cuda", generator=gen)
    ob = torch.empty_like(xb)
    c.scan_i32(big, xb.data_ptr(), ob.data_ptr(), 0, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert torch.equal(ob[1:] - ob[:-1], xb[1:]) and int(ob[0]) == int(xb[0])
    for rep in range(3):
        h = dg.int_fill(n, -3, 3, np.int32, tag=900 + rep)
        x.copy_(torch.from_numpy(h))
        torch.cuda.synchronize()
        g.replay()
        torch.cuda.synchronize()
        assert np.array_equal(host(out), oracle.scan_i32(h, seed=3)), rep
    c.close()
