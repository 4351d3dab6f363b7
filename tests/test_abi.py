"""CPU-side checks of the drop-in boundary: libcuxb.so loads and exports exactly what
include/cuxb.h declares (no compute calls -- there is no GPU here)."""
import ctypes
import os
import re

import pytest

from hpc_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "cuxb.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(cuxb_[a-z0-9_]+)\s*\(", src)))


def test_library_is_built_and_loads():
    assert os.path.exists(capi.LIB_PATH), "run python -m hpc_b200.build"
    capi.lib()


def test_every_declared_symbol_is_exported():
    L = ctypes.CDLL(capi.LIB_PATH)
    names = header_functions()
    assert len(names) >= 40
    for n in names:
        assert hasattr(L, n), "declared in cuxb.h but not exported: " + n


def test_python_binding_covers_the_header():
    assert sorted(capi.SIGNATURES) == header_functions()


def test_every_entry_point_cites_the_reference():
    # each block of declarations in the header names the reference file:line it replaces
    src = open(os.path.join(ROOT, "include", "cuxb.h")).read()
    for needle in ("kernel_interface.h:30-35", "kernel_interface.h:61-72", "dot_product.cu", "gemm.cu:266-269",
                   "linear.cpp:199-237", "vector_scan.cpp:21-79", "reduce_fp32.cu:33-116", "operator.cpp:11-23",
                   "array.h:84-114", "timer.h:16-38"):
        assert needle in src, needle


def test_no_gpu_means_a_loud_error_not_a_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(capi.CuxbError):
        capi.Context(0)
    assert capi.lib().cuxb_last_error_string()


def test_version_and_launch_counter():
    L = capi.lib()
    assert b"sm_100a" in L.cuxb_version()
    assert L.cuxb_launch_count() >= 0


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "hpc_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "import oracle" not in text and "from oracle" not in text, f
                assert "liboracle" not in text and "libcuxref" not in text, f
