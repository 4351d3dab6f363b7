"""GPU (needs >= 2 devices, skipped otherwise): the reductions / scan whose cross-GPU exchange is fused into
the kernel (cuxb_mg_*: partials written into every rank's buffer over NVLink peer memory by the finishing
CTA).  Single process, two devices, peer access -- each device's kernel waits for the other's partial, both
are launched asynchronously.  Results must equal the single-device oracle on the concatenated data."""
import numpy as np
import pytest
import torch

import oracle
from hpc_b200 import capi
from hpc_b200 import datagen as dg
from hpc_b200 import multi

pytestmark = pytest.mark.gpu


def _setup(world):
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    L = capi.lib()
    for a in range(world):
        for b in range(world):
            if a != b:
                capi.check(L.cuxb_enable_peer_access(a, b), "cuxb_enable_peer_access")
    ctxs = [capi.Context(r) for r in range(world)]
    bufs = [torch.zeros(64, dtype=torch.int64, device="cuda:%d" % r) for r in range(world)]
    # Load every kernel on every device BEFORE kernels start waiting on each other: the first launch of
    # a kernel on a device loads its module, and with peer access enabled that can wait for the other
    # device to go idle -- which a kernel spinning on a peer's partial never does (single process only).
    for r in range(world):
        x = torch.ones(4096, device="cuda:%d" % r)
        xi = torch.ones(4096, dtype=torch.int32, device="cuda:%d" % r)
        o = torch.zeros(4096, dtype=torch.int32, device="cuda:%d" % r)
        c = ctxs[r]
        c.dot_f32(4096, x.data_ptr(), x.data_ptr(), x.data_ptr())
        c.sum_f32(4096, x.data_ptr(), x.data_ptr())
        c.nrm2_f32(4096, x.data_ptr(), x.data_ptr())
        c.sum_i32(4096, xi.data_ptr(), o.data_ptr())
        c.scan_i32(4096, xi.data_ptr(), o.data_ptr(), 0)
        torch.cuda.synchronize(r)
    return ctxs, bufs, [b.data_ptr() for b in bufs]


@pytest.mark.parametrize("world", [2])
def test_fused_exchange_reductions_and_scan(world):
    ctxs, bufs, ptrs = _setup(world)
    n = 3_000_001
    a, b = dg.int_fill(n, -2, 2, tag=1), dg.int_fill(n, -2, 2, tag=2)
    x = np.full(n, 3000, np.int32)
    x[::5] = -7
    spans = [multi.shard_range(n, world, r) for r in range(world)]
    da = [torch.from_numpy(a[lo:hi]).to("cuda:%d" % r) for r, (lo, hi) in enumerate(spans)]
    db = [torch.from_numpy(b[lo:hi]).to("cuda:%d" % r) for r, (lo, hi) in enumerate(spans)]
    dx = [torch.from_numpy(x[lo:hi]).to("cuda:%d" % r) for r, (lo, hi) in enumerate(spans)]
    res = [torch.full((1,), -1.0, device="cuda:%d" % r) for r in range(world)]
    out = [torch.empty_like(t) for t in dx]
    for rep in range(3):  # repeated calls: epochs / slot parity
        for r in range(world):
            ctxs[r].mg_dot_f32(da[r].numel(), da[r].data_ptr(), db[r].data_ptr(), res[r].data_ptr(), world, r, ptrs)
        for r in range(world):
            torch.cuda.synchronize(r)
        for r in range(world):
            assert float(res[r][0]) == float(oracle.dot_f32(a, b))  # identical on every rank, exact (integer fill)
    for r in range(world):
        ctxs[r].mg_sum_f32(da[r].numel(), da[r].data_ptr(), res[r].data_ptr(), world, r, ptrs)
    for r in range(world):
        torch.cuda.synchronize(r)
        assert float(res[r][0]) == float(oracle.sum_f32(a))
    for r in range(world):
        ctxs[r].mg_nrm2_f32(da[r].numel(), da[r].data_ptr(), res[r].data_ptr(), world, r, ptrs)
    for r in range(world):
        torch.cuda.synchronize(r)
        assert float(res[r][0]) == float(oracle.nrm2_f32(a))
    want = oracle.scan_i32(x)
    assert want.min() < 0
    for rep in range(2):
        for r in range(world):
            ctxs[r].mg_scan_i32(dx[r].numel(), dx[r].data_ptr(), out[r].data_ptr(), world, r, ptrs)
        for r in range(world):
            torch.cuda.synchronize(r)
        got = np.concatenate([o.cpu().numpy() for o in out])
        assert np.array_equal(got, want)  # bit-exact across the shard boundary, wrap-around included


def test_mg_argument_errors():
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    c = capi.Context(0)
    t = torch.zeros(16, device="cuda:0")
    with pytest.raises(capi.CuxbError):
        c.mg_dot_f32(16, t.data_ptr(), t.data_ptr(), t.data_ptr(), 9, 0, [0] * 9)  # world > 8
    with pytest.raises(capi.CuxbError):
        c.mg_dot_f32(16, t.data_ptr(), t.data_ptr(), t.data_ptr(), 2, 0, [t.data_ptr(), 0])  # null peer buffer
    # world == 1 degenerates to the single-GPU kernel
    c.mg_sum_f32(16, t.data_ptr(), t.data_ptr(), 1, 0, [t.data_ptr()])
    torch.cuda.synchronize()


def test_fused_b_broadcast_gemm_two_gpus():
    """Row-sharded fp16 GEMM: rank 0 owns B, rank 1's kernel pulls B over NVLink while it multiplies."""
    world = 2
    ctxs, _, _ = _setup(world)
    M, N, K = 1024, 2304, 1000  # ragged: N not a multiple of 256, K not a multiple of 24 or 64
    A = [oracle.f32_to_f16(dg.int_fill((M, K), -10, 9, tag=90 + r)) for r in range(world)]
    B16 = oracle.f32_to_f16(dg.int_fill((K, N), -10, 9, tag=95))
    dev = lambda x, r: torch.from_numpy(np.ascontiguousarray(x).view(np.int16)).to("cuda:%d" % r)
    dA = [dev(A[r], r) for r in range(world)]
    dB_root = dev(B16, 0)
    dB_local = torch.zeros((K, N), dtype=torch.int16, device="cuda:1")
    dC = [torch.zeros((M, N), dtype=torch.int16, device="cuda:%d" % r) for r in range(world)]
    # load the GEMM kernel on both devices before anything waits on a peer (see _setup)
    for r in range(world):
        ctxs[r].hgemm_f16(M, N, K, 1.0, dA[r].data_ptr(), K, (dB_root if r == 0 else dB_local).data_ptr(), N, 0.0,
                          dC[r].data_ptr(), N, capi.FLOAT16)
        torch.cuda.synchronize(r)
    for rep in range(3):
        dB_local.zero_()
        torch.cuda.synchronize(1)
        ctxs[0].mg_hgemm_f16(M, N, K, 1.0, dA[0].data_ptr(), K, dB_root.data_ptr(), 0, N, 0.0, dC[0].data_ptr(), N, True)
        ctxs[1].mg_hgemm_f16(M, N, K, 1.0, dA[1].data_ptr(), K, dB_root.data_ptr(), dB_local.data_ptr(), N, 0.0,
                             dC[1].data_ptr(), N, False)
        for r in range(world):
            torch.cuda.synchronize(r)
        assert np.array_equal(dB_local.cpu().numpy().view(np.uint16), B16)  # the broadcast itself
        for r in range(world):
            want = oracle.hgemm_f16(A[r], B16, threads=8)
            assert np.array_equal(dC[r].cpu().numpy().view(np.uint16), want), "rank %d rep %d" % (r, rep)


def test_fused_b_broadcast_sgemm_and_int8_two_gpus():
    """Row-sharded fp32 SGEMM and int8 linear: rank 1's launch copies B from rank 0 over NVLink (copy CTAs /
    the idle warp) while its other CTAs multiply the panels that have arrived."""
    world = 2
    ctxs, _, _ = _setup(world)
    # ---- fp32
    M, N, K = 1024, 1152, 528
    A = [dg.int_fill((M, K), -2, 5, tag=70 + r) for r in range(world)]
    B = dg.int_fill((K, N), -2, 5, tag=75)
    dA = [torch.from_numpy(A[r]).to("cuda:%d" % r) for r in range(world)]
    dB_root = torch.from_numpy(B).to("cuda:0")
    dB_local = torch.zeros((K, N), device="cuda:1")
    dC = [torch.zeros((M, N), device="cuda:%d" % r) for r in range(world)]
    for r in range(world):
        torch.cuda.synchronize(r)
    for rep in range(3):
        dB_local.zero_()
        torch.cuda.synchronize(1)
        ctxs[0].mg_sgemm_f32(M, N, K, 1.0, dA[0].data_ptr(), K, dB_root.data_ptr(), 0, N, 0.0, dC[0].data_ptr(), N, True)
        ctxs[1].mg_sgemm_f32(M, N, K, 1.0, dA[1].data_ptr(), K, dB_root.data_ptr(), dB_local.data_ptr(), N, 0.0,
                             dC[1].data_ptr(), N, False)
        for r in range(world):
            torch.cuda.synchronize(r)
        assert np.array_equal(dB_local.cpu().numpy(), B)
        for r in range(world):
            want = oracle.gemm_f32(A[r], B, np.zeros((M, N), np.float32), 1.0, 0.0, order="ikj", threads=8)
            assert np.array_equal(dC[r].cpu().numpy().view(np.uint32), want.view(np.uint32)), "rank %d rep %d" % (r, rep)
    # ---- int8
    M, N, K = 512, 1100, 2064
    A8 = [dg.int_fill((M, K), -128, 127, tag=80 + r).astype(np.int8) for r in range(world)]
    Bt = dg.int_fill((N, K), -128, 127, tag=85).astype(np.int8)
    dA8 = [torch.from_numpy(A8[r]).to("cuda:%d" % r) for r in range(world)]
    dBt_root = torch.from_numpy(Bt).to("cuda:0")
    dBt_local = torch.zeros((N, K), dtype=torch.int8, device="cuda:1")
    dC8 = [torch.zeros((M, N), dtype=torch.int32, device="cuda:%d" % r) for r in range(world)]
    for rep in range(3):
        dBt_local.zero_()
        torch.cuda.synchronize(1)
        ctxs[0].mg_linear_s8(M, N, K, dA8[0].data_ptr(), dBt_root.data_ptr(), 0, dC8[0].data_ptr(), True)
        ctxs[1].mg_linear_s8(M, N, K, dA8[1].data_ptr(), dBt_root.data_ptr(), dBt_local.data_ptr(), dC8[1].data_ptr(), False)
        for r in range(world):
            torch.cuda.synchronize(r)
        assert np.array_equal(dBt_local.cpu().numpy(), Bt)
        for r in range(world):
            assert np.array_equal(dC8[r].cpu().numpy(), oracle.linear_s8(A8[r], Bt, threads=8)), "rank %d rep %d" % (r, rep)


def test_chained_broadcast_across_two_gpus():
    """cuxb_mg_set_chain with links that really wait: the root's B is on GPU 0, 'rank 1' (GPU 1) pulls it,
    'rank 2' (a second context on GPU 0) pulls rank 1's copy from GPU 1 panel by panel as it lands there."""
    world = 2
    ctxs, _, _ = _setup(world)
    c1, c2 = ctxs[1], capi.Context(0)
    nflag = capi.MG_CHAIN_FLAGS_BYTES // 4
    f1 = torch.zeros(nflag, dtype=torch.int32, device="cuda:1")
    f2 = torch.zeros(nflag, dtype=torch.int32, device="cuda:0")
    g0 = torch.Generator(device="cuda:0")
    g0.manual_seed(3)
    for kind in ("f16", "f32"):
        if kind == "f16":
            M, N, K = 1024, 2304, 2000
            mk = lambda shape, lo, hi: torch.randint(lo, hi, shape, device="cuda:0", generator=g0).half()
            A0, B0 = mk((M, K), -10, 10), mk((K, N), -10, 10)
            plain = lambda c, A, B, C: c.hgemm_f16(M, N, K, 1.0, A.data_ptr(), K, B.data_ptr(), N, 0.0, C.data_ptr(), N, capi.FLOAT16)
            mg = lambda c, A, Bup, Bl, C: c.mg_hgemm_f16(M, N, K, 1.0, A.data_ptr(), K, Bup.data_ptr(), Bl.data_ptr(), N, 0.0,
                                                         C.data_ptr(), N, False)
        else:
            M, N, K = 1024, 1152, 528
            mk = lambda shape, lo, hi: torch.randint(lo, hi, shape, device="cuda:0", generator=g0).float()
            A0, B0 = mk((M, K), -2, 6), mk((K, N), -2, 6)
            plain = lambda c, A, B, C: c.sgemm_f32(M, N, K, 1.0, A.data_ptr(), K, B.data_ptr(), N, 0.0, C.data_ptr(), N)
            mg = lambda c, A, Bup, Bl, C: c.mg_sgemm_f32(M, N, K, 1.0, A.data_ptr(), K, Bup.data_ptr(), Bl.data_ptr(), N, 0.0,
                                                         C.data_ptr(), N, False)
        A1 = A0.to("cuda:1")
        want = torch.zeros((M, N), dtype=A0.dtype, device="cuda:0")
        plain(ctxs[0], A0, B0, want)
        B1, C1 = torch.zeros_like(B0, device="cuda:1"), torch.zeros_like(want, device="cuda:1")
        B2, C2 = torch.zeros_like(B0), torch.zeros_like(want)
        # load the pulling instantiations on both devices before any launch waits on the other GPU
        mg(c1, A1, B1, B1.clone(), C1)
        mg(c2, A0, B0, B2, C2)
        for r in range(world):
            torch.cuda.synchronize(r)
        c1.mg_set_chain(f1.data_ptr(), 0, f2.data_ptr())
        c2.mg_set_chain(f2.data_ptr(), f1.data_ptr(), 0)
        for rep in range(3):
            B1.zero_(), B2.zero_(), C1.zero_(), C2.zero_()
            for r in range(world):
                torch.cuda.synchronize(r)
            mg(c1, A1, B0, B1, C1)     # GPU 1 pulls from the root on GPU 0
            mg(c2, A0, B1, B2, C2)     # GPU 0 pulls from GPU 1's copy while it is being filled
            for r in range(world):
                torch.cuda.synchronize(r)
            assert torch.equal(B1.cpu(), B0.cpu()) and torch.equal(B2, B0), (kind, rep)
            assert torch.equal(C1.cpu(), want.cpu()) and torch.equal(C2, want), (kind, rep)
        c1.mg_set_chain(0, 0, 0)
        c2.mg_set_chain(0, 0, 0)


def test_fused_exchange_replayed_from_cuda_graphs():
    """One graph per GPU, each holding that rank's mg_dot + mg_scan; replayed together on fresh data.  The
    exchange epoch is counted in the exchange buffer, so every replay runs under a new one."""
    world = 2
    ctxs, bufs, ptrs = _setup(world)
    n = 500_003
    da = [torch.zeros(n, device="cuda:%d" % r) for r in range(world)]
    db = [torch.zeros(n, device="cuda:%d" % r) for r in range(world)]
    dx = [torch.zeros(n, dtype=torch.int32, device="cuda:%d" % r) for r in range(world)]
    res = [torch.zeros(1, device="cuda:%d" % r) for r in range(world)]
    out = [torch.zeros_like(t) for t in dx]
    streams = [torch.cuda.Stream(device=r) for r in range(world)]

    def work(r, st):
        ctxs[r].mg_dot_f32(n, da[r].data_ptr(), db[r].data_ptr(), res[r].data_ptr(), world, r, ptrs, st)
        ctxs[r].mg_scan_i32(n, dx[r].data_ptr(), out[r].data_ptr(), world, r, ptrs, st)
    # warm-up: both ranks must run it together (they wait for each other)
    for r in range(world):
        with torch.cuda.device(r):
            work(r, streams[r].cuda_stream)
    for r in range(world):
        torch.cuda.synchronize(r)
    graphs = []
    for r in range(world):
        with torch.cuda.device(r):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=streams[r]):
                work(r, torch.cuda.current_stream().cuda_stream)
            graphs.append(g)
    for rep in range(3):
        a = [dg.int_fill(n, -2, 2, tag=300 + 10 * rep + r) for r in range(world)]
        b = [dg.int_fill(n, -2, 2, tag=350 + 10 * rep + r) for r in range(world)]
        x = [dg.int_fill(n, -5, 5, np.int32, tag=400 + 10 * rep + r) for r in range(world)]
        for r in range(world):
            da[r].copy_(torch.from_numpy(a[r])), db[r].copy_(torch.from_numpy(b[r])), dx[r].copy_(torch.from_numpy(x[r]))
            torch.cuda.synchronize(r)
        for r in range(world):
            with torch.cuda.device(r):
                graphs[r].replay()
        for r in range(world):
            torch.cuda.synchronize(r)
        want_dot = oracle.dot_f32(np.concatenate(a), np.concatenate(b))
        want_scan = oracle.scan_i32(np.concatenate(x))
        for r in range(world):
            assert float(res[r][0]) == float(want_dot), (rep, r)
        assert np.array_equal(np.concatenate([o.cpu().numpy() for o in out]), want_scan), rep


def test_mg_driver_two_gpus_all_ops():
    """cuxb_mg_create: the C-ABI driver that owns contexts, streams, peer access, exchange buffers, flags and
    the copies of B -- the flow of mpi/alg_matrix_multiply.cpp:52-80 (row blocks of A and C, one B, results
    gathered and checked) from one host thread, nothing but C calls."""
    world = 2
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    mg = capi.MultiGpu(range(world))
    dev = lambda x, r: torch.from_numpy(np.ascontiguousarray(x)).to("cuda:%d" % r)
    # reductions + scan over an uneven split
    n = 2_000_003
    a, b = dg.int_fill(n, -2, 2, tag=501), dg.int_fill(n, -2, 2, tag=502)
    x = dg.int_fill(n, -3, 3, np.int32, tag=503)
    spans = [multi.shard_range(n, world, r) for r in range(world)]
    da = [dev(a[lo:hi], r) for r, (lo, hi) in enumerate(spans)]
    db = [dev(b[lo:hi], r) for r, (lo, hi) in enumerate(spans)]
    dx = [dev(x[lo:hi], r) for r, (lo, hi) in enumerate(spans)]
    res = [torch.zeros(1, device="cuda:%d" % r) for r in range(world)]
    out = [torch.empty_like(t) for t in dx]
    for r in range(world):
        torch.cuda.synchronize(r)
    lens = [hi - lo for lo, hi in spans]
    P = lambda ts: [t.data_ptr() for t in ts]
    for rep in range(2):
        mg.all_dot_f32(lens, P(da), P(db), P(res))
        mg.sync()
        assert all(float(t[0]) == float(oracle.dot_f32(a, b)) for t in res)
    mg.all_sum_f32(lens, P(da), P(res))
    mg.sync()
    assert all(float(t[0]) == float(oracle.sum_f32(a)) for t in res)
    mg.all_nrm2_f32(lens, P(da), P(res))
    mg.sync()
    assert all(float(t[0]) == float(oracle.nrm2_f32(a)) for t in res)
    mg.all_scan_i32(lens, P(dx), P(out))
    mg.sync()
    assert np.array_equal(np.concatenate([o.cpu().numpy() for o in out]), oracle.scan_i32(x))
    # row-sharded GEMMs, B on rank 0 only
    M, N, K = 256, 512, 384
    A = [dg.int_fill((M, K), -2, 5, tag=510 + r) for r in range(world)]
    B = dg.int_fill((K, N), -2, 5, tag=515)
    dA, dB = [dev(A[r], r) for r in range(world)], dev(B, 0)
    dC = [torch.zeros((M, N), device="cuda:%d" % r) for r in range(world)]
    for r in range(world):
        torch.cuda.synchronize(r)
    for rep in range(2):
        mg.all_sgemm_f32([M] * world, N, K, 1.0, P(dA), K, dB.data_ptr(), N, 0.0, P(dC), N)
        mg.sync()
        for r in range(world):
            want = oracle.gemm_f32(A[r], B, np.zeros((M, N), np.float32), 1.0, 0.0, order="ikj", threads=8)
            assert np.array_equal(dC[r].cpu().numpy(), want), (r, rep)
    A16 = [oracle.f32_to_f16(dg.int_fill((M, K), -10, 9, tag=520 + r)) for r in range(world)]
    B16 = oracle.f32_to_f16(dg.int_fill((K, N), -10, 9, tag=525))
    dA16 = [dev(A16[r].view(np.int16), r) for r in range(world)]
    dB16 = dev(B16.view(np.int16), 0)
    dC16 = [torch.zeros((M, N), dtype=torch.int16, device="cuda:%d" % r) for r in range(world)]
    for r in range(world):
        torch.cuda.synchronize(r)
    mg.all_hgemm_f16([M] * world, N, K, 1.0, P(dA16), K, dB16.data_ptr(), N, 0.0, P(dC16), N, capi.FLOAT16)
    mg.sync()
    for r in range(world):
        assert np.array_equal(dC16[r].cpu().numpy().view(np.uint16), oracle.hgemm_f16(A16[r], B16, threads=8)), r
    A8 = [dg.int_fill((M, K), -128, 127, tag=530 + r).astype(np.int8) for r in range(world)]
    Bt = dg.int_fill((N, K), -128, 127, tag=535).astype(np.int8)
    dA8, dBt = [dev(A8[r], r) for r in range(world)], dev(Bt, 0)
    dC8 = [torch.zeros((M, N), dtype=torch.int32, device="cuda:%d" % r) for r in range(world)]
    for r in range(world):
        torch.cuda.synchronize(r)
    mg.all_linear_s8([M] * world, N, K, P(dA8), dBt.data_ptr(), P(dC8))
    mg.sync()
    for r in range(world):
        assert np.array_equal(dC8[r].cpu().numpy(), oracle.linear_s8(A8[r], Bt, threads=8)), r
    mg.close()


def test_mg_driver_single_gpu_and_argument_errors():
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    with pytest.raises(capi.CuxbError):
        capi.MultiGpu([0, 0])          # one rank per GPU
    with pytest.raises(capi.CuxbError):
        capi.MultiGpu([99])
    mg = capi.MultiGpu([0])            # a world of one degenerates to the single-GPU kernels
    x = torch.ones(1000, device="cuda:0")
    r = torch.zeros(1, device="cuda:0")
    mg.all_sum_f32([1000], [x.data_ptr()], [r.data_ptr()])
    mg.sync()
    assert float(r[0]) == 1000.0
    mg.close()
