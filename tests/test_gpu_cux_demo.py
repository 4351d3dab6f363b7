"""GPU: the cux demo (hpc_b200/cux/example/demo.cpp) -- cux's own DotProductTest / Nrm2Test / GEMMTest
flow (0-frameworks/cux/example/demo.cpp) plus the three added ops -- through Executor / OpFactory /
plugins / Array4D over the C ABI.  Every ResultChecker verdict must be Pass and host kernel 0 must
equal device kernel 0 bit for bit (integer-valued fills)."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_demo(*args):
    # use the binaries built by __graft_entry__.build(); never rebuild libcuxb.so here -- this process
    # has it loaded (other tests), and replacing a mapped library under a running process is unsafe
    demo = os.path.join(ROOT, "hpc_b200", "cux", "cux_demo")
    if not os.path.exists(demo):
        from hpc_b200.cux import build as hb
        _, demo, _ = hb.build()
    return subprocess.run([demo, *args], capture_output=True, text=True, timeout=250)


def test_demo_reference_shapes():
    r = run_demo()  # 4 096 000-element vectors, gemm 512x480x576 alpha=1 beta=3
    out = r.stdout
    assert r.returncode == 0, out[-3000:] + r.stderr[-2000:]
    assert "cux_demo: all checks passed" in out
    assert "Fail" not in out
    for name in ("dot", "nrm2", "gemm fp32", "reduce_sum", "linear_int8", "linear_bf16", "scan_i32"):
        assert "host V0 vs device V0 (%s): 0 of" % name in out, name
    assert out.count("Pass: V0 vs V") >= 3  # plugin kernels and the fp16 tcgen05 kernel vs device V0
    assert "tcgen05" in out and "Plugin example" in out


def test_demo_quick():
    r = run_demo("--quick")
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]


def test_demo_with_device_side_fill_and_pinned_host_memory():
    # Executor::BindAndFillOnDevice (fill + dtype conversion on the GPU) and pinned Array4D host copies
    r = run_demo("--quick", "--device-fill", "--pinned")
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert "cux_demo: all checks passed" in r.stdout and "Fail" not in r.stdout


def test_reference_data_test_cases_on_the_device():
    """hpc_b200/cux/tests/device_test.cpp: the reference's own DataTest suite (tests/data_test.cpp:104-143 --
    push and back for int32 / float / half / int8, conversions, Save / Restore) against this Array4D, with
    pageable and pinned host copies, plus the device-side extensions."""
    exe = os.path.join(ROOT, "hpc_b200", "cux", "cux_device_test")
    assert os.path.exists(exe), "run __graft_entry__.build() first"
    r = subprocess.run([exe], capture_output=True, text=True, timeout=200)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert "cux_device_test: all passed" in r.stdout and "FAILED" not in r.stdout


def _gpu_count():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("gpus", [1, 2])
def test_demo_multi_gpu_executor(gpus):
    """cux::MultiExecutor (Executor::Initialize generalised to g devices over the C ABI's cuxb_mg driver): dot and
    scan over contiguous slices, gemm over row blocks with B on rank 0 only -- mpi/alg_matrix_multiply.cpp:52-80's
    flow -- checked by cux's ResultChecker against the host kernels.  One GPU is the degenerate world of one."""
    if _gpu_count() < gpus:
        pytest.skip("needs %d GPUs" % gpus)
    r = run_demo("--quick", "--multi-only", "--gpus", str(gpus))
    out = r.stdout
    assert r.returncode == 0, out[-3000:] + r.stderr[-2000:]
    assert "cux_demo: all checks passed" in out and "FAIL" not in out and "Fail" not in out
    for what in ("multi-GPU dot", "multi-GPU gemm fp32", "multi-GPU scan_i32"):
        assert what in out, what
