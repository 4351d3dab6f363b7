"""GPU (one device is enough): the GEMM kernels whose broadcast of B is fused into the launch
(cuxb_mg_sgemm_f32, cuxb_mg_linear_s8_s32, cuxb_mg_hgemm_f16_f32acc).  On a non-root rank part of the
launch copies B from the root's memory while the rest multiplies the panels that have arrived.  The copy
does not care where the source lives, so here the "root's B" is simply another buffer on the same GPU:
the panel flags, the role tickets and the arithmetic are exactly those of the multi-GPU run
(tests/test_gpu_multi.py repeats it across two devices over NVLink).

Bar: the local copy of B equals the root's byte for byte, C is BIT-EXACT against the oracle / the
single-GPU kernel, on repeated calls (the flags only grow) and on shapes with more CTAs than the GPU holds."""
import numpy as np
import pytest
import torch

import oracle
from hpc_b200 import capi
from hpc_b200 import datagen as dg
from gpu_util import bits32, ctx, dev, host, sync

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape,pad", [((256, 384, 400), 0), ((128, 128, 16), 0), ((384, 256, 272), 8)])
def test_sgemm_pull_small_vs_oracle(shape, pad):
    M, N, K = shape
    ldb = N + pad
    A, B, C = dg.int_fill((M, K), -2, 5, tag=11), dg.int_fill((K, N), -2, 5, tag=12), dg.int_fill((M, N), -2, 5, tag=13)
    want = oracle.gemm_f32(A, B, C, 1.0, 3.0, order="ikj", threads=8)
    Bp = np.full((K, ldb), 77.0, np.float32)
    Bp[:, :N] = B
    dA, dBroot = dev(A), dev(Bp)
    for rep in range(3):
        dBloc = torch.full((K, ldb), -5.0, device="cuda")
        dC = dev(C)
        sync()
        ctx().mg_sgemm_f32(M, N, K, 1.0, dA.data_ptr(), K, dBroot.data_ptr(), dBloc.data_ptr(), ldb, 3.0,
                           dC.data_ptr(), N, False)
        sync()
        got_b = host(dBloc)
        assert np.array_equal(got_b[:, :N], B), "rep %d" % rep
        if pad:
            assert np.all(got_b[:, N:] == -5.0)          # the padding columns are not touched
        assert np.array_equal(bits32(host(dC)), bits32(want)), "rep %d" % rep


def test_sgemm_pull_many_waves_vs_single_gpu_kernel():
    M, N, K = 2560, 2048, 1040   # 320 tiles + 64 copy CTAs > 296 resident; K = 8 panels + a 16-row tail
    g = torch.Generator(device="cuda")
    g.manual_seed(5)
    A = torch.randint(-2, 6, (M, K), device="cuda", generator=g).float()
    B = torch.randint(-2, 6, (K, N), device="cuda", generator=g).float()
    C0 = torch.randint(-2, 6, (M, N), device="cuda", generator=g).float()
    want = C0.clone()
    ctx().sgemm_f32(M, N, K, 2.0, A.data_ptr(), K, B.data_ptr(), N, -1.0, want.data_ptr(), N)
    for rep in range(2):
        Bl = torch.zeros_like(B)
        got = C0.clone()
        sync()
        ctx().mg_sgemm_f32(M, N, K, 2.0, A.data_ptr(), K, B.data_ptr(), Bl.data_ptr(), N, -1.0, got.data_ptr(), N, False)
        sync()
        assert torch.equal(Bl, B)
        assert torch.equal(got, want)
    # the root's call is the plain kernel
    got = C0.clone()
    ctx().mg_sgemm_f32(M, N, K, 2.0, A.data_ptr(), K, B.data_ptr(), 0, N, -1.0, got.data_ptr(), N, True)
    sync()
    assert torch.equal(got, want)


def test_sgemm_pull_rejects_ragged_shapes():
    t = torch.zeros(256 * 256, device="cuda")
    with pytest.raises(capi.CuxbError, match="128"):
        ctx().mg_sgemm_f32(100, 128, 16, 1.0, t.data_ptr(), 16, t.data_ptr(), t.data_ptr(), 128, 0.0, t.data_ptr(), 128,
                           False)
    with pytest.raises(capi.CuxbError):
        ctx().mg_sgemm_f32(128, 128, 16, 1.0, t.data_ptr(), 16, 0, t.data_ptr(), 128, 0.0, t.data_ptr(), 128, False)


@pytest.mark.parametrize("shape", [(512, 600, 1008), (256, 256, 256), (128, 4, 16), (300, 2304, 4096)])
def test_linear_s8_pull_bit_exact(shape):
    M, N, K = shape
    A = dg.int_fill((M, K), -128, 127, tag=21).astype(np.int8)
    Bt = dg.int_fill((N, K), -128, 127, tag=22).astype(np.int8)
    want = oracle.linear_s8(A, Bt, threads=8)
    dA, dBroot = dev(A), dev(Bt)
    for rep in range(3):
        dBloc = torch.full((N, K), 3, dtype=torch.int8, device="cuda")
        dC = torch.full((M, N), -1, dtype=torch.int32, device="cuda")
        sync()
        ctx().mg_linear_s8(M, N, K, dA.data_ptr(), dBroot.data_ptr(), dBloc.data_ptr(), dC.data_ptr(), False)
        sync()
        assert np.array_equal(host(dBloc), Bt), "rep %d" % rep
        assert np.array_equal(host(dC), want), "rep %d" % rep
    dC = torch.full((M, N), -1, dtype=torch.int32, device="cuda")
    ctx().mg_linear_s8(M, N, K, dA.data_ptr(), dBroot.data_ptr(), 0, dC.data_ptr(), True)   # root: in place
    sync()
    assert np.array_equal(host(dC), want)


def test_hgemm_pull_same_device():
    M, N, K = 512, 1280, 1000
    A = oracle.f32_to_f16(dg.int_fill((M, K), -10, 9, tag=31))
    B = oracle.f32_to_f16(dg.int_fill((K, N), -10, 9, tag=32))
    want = oracle.hgemm_f16(A, B, threads=8)
    dA, dBroot = dev(A), dev(B)
    for rep in range(2):
        dBloc = torch.zeros((K, N), dtype=torch.int16, device="cuda")
        dC = torch.zeros((M, N), dtype=torch.int16, device="cuda")
        sync()
        ctx().mg_hgemm_f16(M, N, K, 1.0, dA.data_ptr(), K, dBroot.data_ptr(), dBloc.data_ptr(), N, 0.0, dC.data_ptr(), N,
                           False)
        sync()
        assert np.array_equal(host(dBloc, np.uint16), B)
        assert np.array_equal(host(dC, np.uint16), want)


def test_pull_flags_survive_a_change_of_panel_count():
    """Same grid (148 CTAs), different number of B panels from one call to the next: the readiness flags
    are per panel and only grow, so the library has to notice and start them over."""
    M, K = 4096, 256
    g = torch.Generator(device="cuda")
    g.manual_seed(9)
    A = torch.randint(-10, 10, (M, K), device="cuda", generator=g).half()
    for N in (1280, 2304, 1280, 768):
        B = torch.randint(-10, 10, (K, N), device="cuda", generator=g).half()
        want = torch.zeros((M, N), dtype=torch.float16, device="cuda")
        ctx().hgemm_f16(M, N, K, 1.0, A.data_ptr(), K, B.data_ptr(), N, 0.0, want.data_ptr(), N, capi.FLOAT16)
        Bl = torch.zeros_like(B)
        got = torch.zeros_like(want)
        sync()
        ctx().mg_hgemm_f16(M, N, K, 1.0, A.data_ptr(), K, B.data_ptr(), Bl.data_ptr(), N, 0.0, got.data_ptr(), N, False)
        sync()
        assert torch.equal(Bl, B) and torch.equal(got, want), N
    Af = A.float()
    for Kf in (256, 128, 256):
        B = torch.randint(-2, 6, (Kf, 1280), device="cuda", generator=g).float()
        want = torch.zeros((M, 1280), device="cuda")
        ctx().sgemm_f32(M, 1280, Kf, 1.0, Af.data_ptr(), K, B.data_ptr(), 1280, 0.0, want.data_ptr(), 1280)
        Bl = torch.zeros_like(B)
        got = torch.zeros_like(want)
        sync()
        ctx().mg_sgemm_f32(M, 1280, Kf, 1.0, Af.data_ptr(), K, B.data_ptr(), Bl.data_ptr(), 1280, 0.0, got.data_ptr(),
                           1280, False)
        sync()
        assert torch.equal(Bl, B) and torch.equal(got, want), Kf


def test_chained_broadcast_flag_protocol_sequential():
    """cuxb_mg_set_chain on one GPU: 'rank 1' pulls from the root's B, 'rank 2' from rank 1's copy.  The
    two launches run one after the other on the default stream, so nothing ever spins here -- this checks the
    epoch arithmetic (panel flags, the call counter, the downstream rank's "I have read it" word)
    over repeated calls and changing shapes; tests/test_gpu_multi.py runs a chain whose links really wait."""
    c1, c2 = capi.Context(0), capi.Context(0)
    nflag = capi.MG_CHAIN_FLAGS_BYTES // 4
    f1 = torch.zeros(nflag, dtype=torch.int32, device="cuda")
    f2 = torch.zeros(nflag, dtype=torch.int32, device="cuda")
    c1.mg_set_chain(f1.data_ptr(), 0, f2.data_ptr())
    c2.mg_set_chain(f2.data_ptr(), f1.data_ptr(), 0)
    g = torch.Generator(device="cuda")
    g.manual_seed(11)
    epoch = 0
    for (M, N, K) in [(512, 1280, 520), (512, 768, 256), (512, 1280, 520)]:
        A = torch.randint(-10, 10, (M, K), device="cuda", generator=g).half()
        for rep in range(2):
            B = torch.randint(-10, 10, (K, N), device="cuda", generator=g).half()
            want = torch.zeros((M, N), dtype=torch.float16, device="cuda")
            ctx().hgemm_f16(M, N, K, 1.0, A.data_ptr(), K, B.data_ptr(), N, 0.0, want.data_ptr(), N, capi.FLOAT16)
            B1, B2 = torch.zeros_like(B), torch.zeros_like(B)
            C1, C2 = torch.zeros_like(want), torch.zeros_like(want)
            sync()
            c1.mg_hgemm_f16(M, N, K, 1.0, A.data_ptr(), K, B.data_ptr(), B1.data_ptr(), N, 0.0, C1.data_ptr(), N, False)
            c2.mg_hgemm_f16(M, N, K, 1.0, A.data_ptr(), K, B1.data_ptr(), B2.data_ptr(), N, 0.0, C2.data_ptr(), N, False)
            sync()
            epoch += 1
            assert torch.equal(B1, B) and torch.equal(B2, B)
            assert torch.equal(C1, want) and torch.equal(C2, want)
            panels = (N + 255) // 256
            # a rank's flags are written by its neighbours: panel words by the rank it pulls from, word 4096 by the
            # rank that pulls from it; word 4097 is its own count of completed calls
            a1, a2 = f1.cpu().numpy(), f2.cpu().numpy()
            assert np.all(a2[:panels] == epoch) and a2[4097] == epoch and a2[4096] == 0      # rank 2: fed by rank 1, nobody downstream
            assert np.all(a1[:panels] == 0) and a1[4096] == epoch and a1[4097] == epoch      # rank 1: fed by the root (no flags)
    # the same flags carry on through the fp32 and int8 kernels (one epoch sequence per chain)
    M, N, K = 384, 256, 272
    A = torch.randint(-2, 6, (M, K), device="cuda", generator=g).float()
    B = torch.randint(-2, 6, (K, N), device="cuda", generator=g).float()
    want = torch.zeros((M, N), device="cuda")
    ctx().sgemm_f32(M, N, K, 1.0, A.data_ptr(), K, B.data_ptr(), N, 0.0, want.data_ptr(), N)
    for rep in range(2):
        B1, B2, C1, C2 = torch.zeros_like(B), torch.zeros_like(B), torch.zeros_like(want), torch.zeros_like(want)
        sync()
        c1.mg_sgemm_f32(M, N, K, 1.0, A.data_ptr(), K, B.data_ptr(), B1.data_ptr(), N, 0.0, C1.data_ptr(), N, False)
        c2.mg_sgemm_f32(M, N, K, 1.0, A.data_ptr(), K, B1.data_ptr(), B2.data_ptr(), N, 0.0, C2.data_ptr(), N, False)
        sync()
        epoch += 1
        assert torch.equal(B2, B) and torch.equal(C1, want) and torch.equal(C2, want)
        assert int(f1[4096]) == epoch and int(f1[4097]) == epoch and int(f2[4097]) == epoch and int(f2[2]) == epoch  # 3 k-panels
    A8 = torch.randint(-128, 128, (256, 512), device="cuda", generator=g, dtype=torch.int32).to(torch.int8)
    Bt = torch.randint(-128, 128, (600, 512), device="cuda", generator=g, dtype=torch.int32).to(torch.int8)
    want = torch.zeros((256, 600), dtype=torch.int32, device="cuda")
    ctx().linear_s8(256, 600, 512, A8.data_ptr(), Bt.data_ptr(), want.data_ptr())
    B1, B2 = torch.zeros_like(Bt), torch.zeros_like(Bt)
    C1, C2 = torch.zeros_like(want), torch.zeros_like(want)
    sync()
    c1.mg_linear_s8(256, 600, 512, A8.data_ptr(), Bt.data_ptr(), B1.data_ptr(), C1.data_ptr(), False)
    c2.mg_linear_s8(256, 600, 512, A8.data_ptr(), B1.data_ptr(), B2.data_ptr(), C2.data_ptr(), False)
    sync()
    assert torch.equal(B2, Bt) and torch.equal(C1, want) and torch.equal(C2, want)
    assert int(f2[4097]) == epoch + 1 and int(f1[4096]) == epoch + 1
    c1.mg_set_chain(0, 0, 0)
    c2.mg_set_chain(0, 0, 0)
    with pytest.raises(capi.CuxbError):
        c1.mg_set_chain(0, f1.data_ptr(), 0)


@pytest.mark.parametrize("kind", ["bf16", "f16"])
def test_linear_16bit_pull_bit_exact(kind):
    """K-major 16-bit Bt pulled in 256-k x 32-row boxes; integer fills, so bit-exact vs the plain kernel and,
    for bf16, vs LinearNormalBf16 (linear.cpp:137-151)."""
    M, N, K = 384, 1096, 1096   # ragged N and K (multiples of 8 only: 16-byte row pitches)
    A = dg.int_fill((M, K), -4, 4, tag=51)
    Bt = dg.int_fill((N, K), -4, 4, tag=52)
    cvt = oracle.f32_to_bf16 if kind == "bf16" else oracle.f32_to_f16
    A16, B16 = cvt(A), cvt(Bt)
    dt = capi.BFLOAT16 if kind == "bf16" else capi.FLOAT16
    dA, dBroot = dev(A16), dev(B16)
    want = torch.zeros((M, N), dtype=torch.int16, device="cuda")
    (ctx().linear_bf16 if kind == "bf16" else ctx().linear_f16)(M, N, K, dA.data_ptr(), dBroot.data_ptr(), want.data_ptr(), dt)
    sync()
    if kind == "bf16":
        assert np.array_equal(host(want, np.uint16), oracle.linear_bf16(A16, B16, threads=8))
    for rep in range(2):
        dBloc = torch.zeros((N, K), dtype=torch.int16, device="cuda")
        dC = torch.zeros((M, N), dtype=torch.int16, device="cuda")
        sync()
        ctx().mg_linear_16(kind, M, N, K, dA.data_ptr(), dBroot.data_ptr(), dBloc.data_ptr(), dC.data_ptr(), dt, False)
        sync()
        assert np.array_equal(host(dBloc, np.uint16), B16)
        assert torch.equal(dC, want)
