"""GPU test of cuxb_scan_i32_noalloc, the scan entry point for callers that cannot allocate.

Who needs it: code that runs where cudaMalloc is off limits -- a plug-in whose host application has
already claimed the GPU's memory for its own pool, a service that must stay inside a fixed memory
budget, or a tight loop that cannot absorb the device-wide stall an allocation causes.  The regular
cuxb_scan_i32 keeps its tile-state scratch in a cuxb_ctx (cudaMalloc); this one uses a fixed array
compiled into libcuxb.so, so it needs neither a context nor any allocation, at the price of a size
limit (2^26 elements) and of all calls on one GPU sharing that array.

The file is named test_zz_* so that pytest collects it after every other test: if it fails under
`pytest -x`, everything else has already run.  It must pass both on its own in a fresh process and
after the rest of the suite.  Integer results are compared bit for bit (the tolerance the other
integer tests use), on two calls in a row that reuse the same scratch.
"""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

import oracle
from hpc_b200 import capi
from hpc_b200 import datagen as dg

pytestmark = pytest.mark.gpu

MAX_ELEMS = 1 << 26


def _noalloc(x, seed):
    out = torch.empty_like(x)
    rc = capi.lib().cuxb_scan_i32_noalloc(torch.cuda.current_stream().cuda_stream, x.numel(), x.data_ptr(),
                                          out.data_ptr(), seed)
    capi.check(rc, "cuxb_scan_i32_noalloc")
    return out


def _with_ctx(ctx, x, seed):
    out = torch.empty_like(x)
    ctx.scan_i32(x.numel(), x.data_ptr(), out.data_ptr(), seed, torch.cuda.current_stream().cuda_stream)
    return out


def test_noalloc_scan_matches_the_allocating_entry_point_twice_in_a_row():
    torch.cuda.set_device(0)
    ctx = capi.Context(0)
    # two calls back to back on the same stream: the second starts from the scratch the first left
    cases = [(5_000_003, 17, 1), ((1 << 22) + 4096, -9, 2), (1, 3, 3), (4096, 0, 4)]
    for n, seed, tag in cases:
        x = torch.from_numpy(dg.int_fill(n, -2, 2, np.int32, tag=tag)).cuda()
        got = _noalloc(x, seed)
        want = _with_ctx(ctx, x, seed)
        torch.cuda.synchronize()
        assert torch.equal(got, want), "n=%d" % n
    # identical input twice: identical output (nothing stale leaks from the shared scratch)
    x = torch.from_numpy(dg.int_fill(3_000_000, -2, 2, np.int32, tag=9)).cuda()
    first, second = _noalloc(x, 5), _noalloc(x, 5)
    torch.cuda.synchronize()
    assert torch.equal(first, second)
    assert torch.equal(first, _with_ctx(ctx, x, 5))
    # wrap-around arithmetic is the same too
    y = torch.full((3_000_000,), 3000, dtype=torch.int32, device="cuda")
    assert torch.equal(_noalloc(y, 0), _with_ctx(ctx, y, 0))
    ctx.close()


def test_noalloc_scan_at_its_size_limit_and_beyond():
    torch.cuda.set_device(0)
    ctx = capi.Context(0)
    g = torch.Generator(device="cuda")
    g.manual_seed(7)
    x = torch.randint(-2, 3, (MAX_ELEMS,), dtype=torch.int32, device="cuda", generator=g)
    assert torch.equal(_noalloc(x, 1), _with_ctx(ctx, x, 1))  # exactly the maximum is accepted
    out = torch.empty(8, dtype=torch.int32, device="cuda")
    rc = capi.lib().cuxb_scan_i32_noalloc(0, MAX_ELEMS + 1, x.data_ptr(), out.data_ptr(), 0)
    assert rc == 100003  # CUXB_ERR_TOO_LARGE
    msg = capi.lib().cuxb_last_error_string().decode()
    assert "67108864" in msg and "cuxb_scan_i32" in msg
    assert capi.lib().cuxb_scan_i32_noalloc(0, -1, x.data_ptr(), out.data_ptr(), 0) != 0
    assert capi.lib().cuxb_scan_i32_noalloc(0, 8, 0, out.data_ptr(), 0) != 0
    assert capi.lib().cuxb_scan_i32_noalloc(0, 0, 0, 0, 0) == 0  # empty input is a no-op
    ctx.close()


def _free_bytes():
    """Free device memory as the driver reports it (cudaMemGetInfo: device-wide, not torch's cache)."""
    return torch.cuda.mem_get_info(0)[0]


def test_noalloc_scan_matches_the_oracle_and_allocates_nothing():
    """The property the entry point exists for: after the first call on a device (which loads the kernel
    and resolves the scratch array's address), calls leave the driver's free-memory figure untouched --
    and the result is the reference recurrence itself, not just "what the other entry point gives"."""
    torch.cuda.set_device(0)
    x_h = dg.int_fill(2_000_003, -2, 2, np.int32, tag=21)
    x = torch.from_numpy(x_h).cuda()
    out = torch.empty_like(x)
    st = torch.cuda.current_stream().cuda_stream
    L = capi.lib()
    capi.check(L.cuxb_scan_i32_noalloc(st, x.numel(), x.data_ptr(), out.data_ptr(), 11), "first call")
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), oracle.scan_i32(x_h, seed=11))
    before = _free_bytes()
    for i in range(100):
        capi.check(L.cuxb_scan_i32_noalloc(st, x.numel() - i, x.data_ptr(), out.data_ptr(), i), "call %d" % i)
    torch.cuda.synchronize()
    assert _free_bytes() == before, "cuxb_scan_i32_noalloc changed the amount of free device memory"
    assert np.array_equal(out.cpu().numpy()[:x.numel() - 99], oracle.scan_i32(x_h[:x.numel() - 99], seed=99))


_FRESH = r"""
import sys
import numpy as np, torch
sys.path.insert(0, %r)
import oracle
from hpc_b200 import capi, datagen as dg
# no capi.Context anywhere in this process: the entry point has to work without one
torch.cuda.set_device(0)
x_h = dg.int_fill(1_000_001, -2, 2, np.int32, tag=33)
x = torch.from_numpy(x_h).cuda()
out = torch.empty_like(x)
for seed in (4, -4):   # two calls in a row
    rc = capi.lib().cuxb_scan_i32_noalloc(torch.cuda.current_stream().cuda_stream, x.numel(), x.data_ptr(), out.data_ptr(), seed)
    assert rc == 0, capi.lib().cuxb_last_error_string()
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), oracle.scan_i32(x_h, seed=seed))
print("fresh-process ok")
"""


def test_noalloc_scan_in_a_process_that_never_creates_a_context():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", _FRESH % root], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "fresh-process ok" in r.stdout, r.stdout + r.stderr
