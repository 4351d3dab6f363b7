"""GPU parity: reductions and prefix scan, called through the C ABI, checked against the oracle.

Bar: integer-valued fp32 fills (the reference's own fill, demo.cpp:30 / data_filler.h:41-48) are
exact in any summation order while |partial| < 2^24 -> BIT-EXACT vs DotHostV0 / ReduceHost /
Nrm2HostV0.  General float fills: |gpu - fp64 truth| <= 4 * eps32 * sum|terms| (a bound the serial
fp32 reference itself is only guaranteed to meet with a factor n), plus cux's own MAE rule.
int32 scan / sum: bit-exact including wrap-around.
"""
import os

import numpy as np
import pytest
import torch

import oracle
from hpc_b200 import datagen as dg
from gpu_util import bits32, ctx, dev, host, sync

pytestmark = pytest.mark.gpu
EPS = np.finfo(np.float32).eps
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_vectors.npz"))

SIZES = [1, 2, 3, 4, 5, 31, 1000, 4099, 262144, 4096000, (1 << 22) + 3]


def run_red(op, a, b=None):
    c = ctx()
    out = torch.full((1,), 123.0, device="cuda")  # must be overwritten, not accumulated
    n = a.numel()
    if op == "dot":
        c.dot_f32(n, a.data_ptr(), b.data_ptr(), out.data_ptr())
    elif op == "sum":
        c.sum_f32(n, a.data_ptr(), out.data_ptr())
    else:
        c.nrm2_f32(n, a.data_ptr(), out.data_ptr())
    sync()
    return host(out)[0]


@pytest.mark.parametrize("n", SIZES)
def test_reductions_integer_fill_bit_exact(n):
    a, b = dg.int_fill(n, -2, 2, tag=n), dg.int_fill(n, -2, 2, tag=n + 1)
    da, db = dev(a), dev(b)
    assert bits32(np.float32(run_red("dot", da, db))) == bits32(np.float32(oracle.dot_f32(a, b)))
    assert bits32(np.float32(run_red("sum", da))) == bits32(np.float32(oracle.sum_f32(a)))
    assert bits32(np.float32(run_red("nrm2", da))) == bits32(np.float32(oracle.nrm2_f32(a)))


@pytest.mark.parametrize("n", [7, 1000, 4096000])
def test_reductions_float_fill_within_bound(n):
    a, b = dg.uniform_fill(n, tag=n), dg.uniform_fill(n, tag=n + 1)
    da, db = dev(a), dev(b)
    truth, abs_sum = oracle.dot_f64(a, b)
    got = run_red("dot", da, db)
    assert abs(got - truth) <= 4 * EPS * abs_sum + 1e-30
    # cux ResultChecker rule (op_assistor.h:38-46): |ref - out| / len < 1e-3 vs DotHostV0
    d, ok = oracle.cux_check(np.array([oracle.dot_f32(a, b)]), np.array([got]), 1.0 / n)
    assert ok, d
    truth, abs_sum = oracle.dot_f64(a)
    assert abs(run_red("sum", da) - truth) <= 4 * EPS * abs_sum
    truth, abs_sum = oracle.dot_f64(a, a)
    assert abs(run_red("nrm2", da) - np.sqrt(truth)) <= 4 * EPS * np.sqrt(truth)


def test_reduction_golden_vectors():
    for name in ("red_int", "red_flt"):
        a, b, w = G[name + "_a"], G[name + "_b"], G[name + "_dot_sum_nrm2"]
        got = np.array([run_red("dot", dev(a), dev(b)), run_red("sum", dev(a)), run_red("nrm2", dev(a))], np.float32)
        if name == "red_int":
            assert np.array_equal(bits32(got), bits32(w))
        else:
            assert np.allclose(got, w, rtol=1e-5)


def test_reduction_edge_cases():
    c = ctx()
    out = torch.full((1,), 9.0, device="cuda")
    x = dev(np.array([-3.5], np.float32))
    c.nrm2_f32(1, x.data_ptr(), out.data_ptr())
    sync()
    assert host(out)[0] == np.float32(3.5)  # Nrm2HostV0: n == 1 -> |x0|
    c.dot_f32(0, 0, 0, out.data_ptr())
    sync()
    assert host(out)[0] == 0.0  # empty input
    # misaligned inputs (not 16-byte aligned) take the scalar path
    a, b = dg.int_fill(10001, -2, 2, tag=5), dg.int_fill(10001, -2, 2, tag=6)
    da, db = dev(a), dev(b)
    c.dot_f32(10000, da.data_ptr() + 4, db.data_ptr() + 4, out.data_ptr())
    sync()
    assert host(out)[0] == oracle.dot_f32(a[1:], b[1:])
    with pytest.raises(Exception):
        c.dot_f32(-1, da.data_ptr(), db.data_ptr(), out.data_ptr())


def test_reduction_is_deterministic_and_repeatable():
    n = 1 << 24
    a, b = dev(dg.uniform_fill(n, tag=1)), dev(dg.uniform_fill(n, tag=2))
    vals = {run_red("dot", a, b).tobytes() for _ in range(5)}
    assert len(vals) == 1


def test_dot_full_size_2p28_bit_exact():
    # BASELINE config 2: N = 2^28 (1 GiB per vector), reference fill [-2,2] -> exact in any order
    n = 1 << 28
    a, b = dg.int_fill(n, -2, 2, tag=281), dg.int_fill(n, -2, 2, tag=282)
    da, db = dev(a), dev(b)
    assert run_red("dot", da, db) == oracle.dot_f32(a, b)
    assert run_red("sum", da) == oracle.sum_f32(a)
    # nrm2: sum(x^2) ~ 5.4e8 > 2^24, so the reference's own serial fp32 accumulation
    # (Nrm2HostV0, nrm2.cpp:11-14) stagnates -- adding 1 stops at 2^24, adding 4 at 2^26 -- and it
    # returns sqrt(2^26) = 8192 instead of ~23170.  Bit parity with a saturated sum is not a goal:
    # the GPU result is checked against the fp64 truth (every x^2 and the total are exact integers).
    truth = float(np.sqrt(np.sum(a.astype(np.float64) ** 2)))
    assert oracle.nrm2_f32(a) == np.float32(8192.0)
    assert abs(run_red("nrm2", da) - truth) <= 2 * EPS * truth


# ------------------------------------------------------------------ scan
def run_scan(x, seed=0, inplace=False):
    c = ctx()
    d = dev(x)
    out = d if inplace else torch.empty_like(d)
    c.scan_i32(d.numel(), d.data_ptr(), out.data_ptr(), seed)
    sync()
    return host(out)


@pytest.mark.parametrize("n", [1, 5, 4095, 4096, 4097, 8191, 8192, 8193, 100000, (1 << 22) + 17, 12345678])
def test_scan_i32_bit_exact(n):
    x = dg.int_fill(n, -2, 2, np.int32, tag=n)
    assert np.array_equal(run_scan(x), oracle.scan_i32(x))
    assert np.array_equal(run_scan(x, seed=-77, inplace=True), oracle.scan_i32(x, seed=-77))


def test_scan_i32_wraparound_and_golden():
    x = np.full(3_000_000, 3000, np.int32)  # crosses 2^31 several times
    want = oracle.scan_i32(x)
    assert want.min() < 0
    assert np.array_equal(run_scan(x), want)
    g = G["scan_in"].astype(np.int32)
    assert np.array_equal(run_scan(g).astype(np.float32), G["scan_out_f32"])
    # misaligned base pointer
    c = ctx()
    d = dev(np.concatenate([[0], g]).astype(np.int32))
    out = torch.empty_like(d)
    c.scan_i32(g.size, d.data_ptr() + 4, out.data_ptr() + 4, 0)
    sync()
    assert np.array_equal(host(out)[1:], oracle.scan_i32(g))


def test_scan_repeated_launches_reuse_tile_state():
    x = dg.int_fill(1 << 20, -2, 2, np.int32, tag=3)
    want = oracle.scan_i32(x)
    for _ in range(4):
        assert np.array_equal(run_scan(x), want)
    y = dg.int_fill(777, -5, 5, np.int32, tag=4)
    assert np.array_equal(run_scan(y), oracle.scan_i32(y))


def test_scan_f32_integer_fill_matches_reference_and_float_tolerance():
    c = ctx()
    x = dg.int_fill(1 << 20, -2, 2, tag=9)
    d = dev(x)
    out = torch.empty_like(d)
    c.scan_f32(x.size, d.data_ptr(), out.data_ptr(), 0.0)
    sync()
    assert np.array_equal(host(out), oracle.scan_f32(x))  # integers: exact in any order
    y = dg.uniform_fill(1 << 18, tag=10)
    d = dev(y)
    out = torch.empty_like(d)
    c.scan_f32(y.size, d.data_ptr(), out.data_ptr(), 0.0)
    sync()
    truth = np.cumsum(y.astype(np.float64))
    bound = 4 * EPS * np.cumsum(np.abs(y).astype(np.float64)) + 1e-6
    assert np.all(np.abs(host(out) - truth) <= bound)


def test_scan_full_size_2p30_properties():
    # BASELINE config 5b: N = 2^30.  Size-independent properties checked on the device:
    # out[i] - out[i-1] == in[i], out[0] == in[0] + seed, and out[-1] == wrap-around sum.
    n = 1 << 30
    g = torch.Generator(device="cuda")
    g.manual_seed(dg.SEED)
    x = torch.randint(-2, 3, (n,), dtype=torch.int32, device="cuda", generator=g)
    out = torch.empty_like(x)
    c = ctx()
    c.scan_i32(n, x.data_ptr(), out.data_ptr(), 11)
    sync()
    assert int(out[0]) == int(x[0]) + 11
    step = 1 << 26
    for s in range(0, n, step):
        e = min(n, s + step)
        lo = max(s, 1)
        assert torch.equal(out[lo:e] - out[lo - 1:e - 1], x[lo:e])
    tot = torch.empty(1, dtype=torch.int32, device="cuda")
    c.sum_i32(n, x.data_ptr(), tot.data_ptr())
    sync()
    assert int(out[-1]) == int(tot[0]) + 11
    # and a CPU cross-check of the first 2^22 elements against the oracle
    m = 1 << 22
    assert np.array_equal(host(out[:m]), oracle.scan_i32(host(x[:m]), seed=11))


def test_sum_i32_wraparound():
    x = np.full(5_000_001, 1000, np.int32)
    c = ctx()
    d = dev(x)
    out = torch.zeros(1, dtype=torch.int32, device="cuda")
    c.sum_i32(x.size, d.data_ptr(), out.data_ptr())
    sync()
    assert host(out)[0] == oracle.sum_i32(x)
