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


def test_host_side_planning_arithmetic_without_a_gpu():
    """The launch planning that is pure host arithmetic, through the library's test hooks: split-K count of the
    SGEMM (sgemm.cu), copy-CTA count of its fused broadcast, chunk size of the host-buffer pipelines
    (host_api.cu).  B200 = 148 SMs -> 296 SGEMM CTA slots, 74 tcgen05 cluster slots."""
    L = capi.lib()
    sm = 148
    # big problems are never split; the reference's split-K case and cux's demo shape are
    assert L.cuxb_debug_sgemm_splits(8192, 8192, 8192, sm) == 1
    assert L.cuxb_debug_sgemm_splits(1280, 2048, 4096, sm) > 1      # 160 tiles on 296 slots
    s = L.cuxb_debug_sgemm_splits(64, 64, 4096, sm)                   # one tile, K = 4096
    assert 16 <= s <= 32
    assert L.cuxb_debug_sgemm_splits(512, 576, 480, sm) in range(2, 8)
    assert L.cuxb_debug_sgemm_splits(128, 128, 128, sm) == 1          # K < 256: not worth a round trip
    for (M, N, K) in [(64, 64, 4096), (512, 576, 480), (256, 256, 1000), (1280, 2048, 4096)]:
        sp = L.cuxb_debug_sgemm_splits(M, N, K, sm)
        tiles = -(-M // 128) * -(-N // 128)
        assert tiles * sp <= 1024 and (sp == 1 or -(-K // sp) >= 64)
    # fused broadcast: the copy CTAs fit into what the last wave of tiles leaves idle
    assert L.cuxb_debug_sgemm_copy_ctas(8192, 8192, sm) == 48          # 4096 tiles = 13 x 296 + 248
    assert L.cuxb_debug_sgemm_copy_ctas(128 * 296, 128, sm) == 64      # no slack at all
    assert L.cuxb_debug_sgemm_copy_ctas(128 * 290, 128, sm) == 64      # 6 idle slots are too few to copy with
    # host pipelines: small problems are one chunk; big ones a multiple of the tile height, at most 16 chunks
    assert L.cuxb_debug_host_chunk_rows(0, 256, 384, 128, sm) == 256
    for kind, (M, N, K), gran in [(0, (8192, 8192, 8192), 128), (1, (16384, 16384, 16384), 256),
                                  (2, (8192, 8192, 8192), 256), (0, (4224, 1024, 1024), 128)]:
        per = L.cuxb_debug_host_chunk_rows(kind, M, N, K, sm)
        assert 0 < per <= M and (per % gran == 0 or per == M)
        assert 2 <= -(-M // per) <= 16


def test_workspace_query_without_a_gpu():
    """cuxb_workspace_bytes: the scratch a call would leave in the context -- pure host arithmetic, ctx may be NULL."""
    import ctypes as C
    L = capi.lib()
    b = C.c_size_t()
    q = lambda k, m, n=0, kk=0: (capi.check(L.cuxb_workspace_bytes(None, k.encode(), m, n, kk, C.byref(b)), "cuxb_workspace_bytes"), b.value)[1]
    assert q("dot_f32", 1 << 28) == 0
    assert q("scan_i32", 1 << 20) == (4096 + 32) * 8                 # the initial capacity covers 2^25 elements
    assert q("scan_i32", 1 << 30) == ((1 << 17) + 32) * 8            # 2^30 / 8192 tiles
    assert q("sgemm_f32", 8192, 8192, 8192) == 0                     # enough tiles: no split-K
    sp = L.cuxb_debug_sgemm_splits(512, 576, 480, 148)
    assert sp > 1 and q("sgemm_f32", 512, 576, 480) == 4096 + 4 * 5 * sp * 16 * 256 * 16
    assert q("hgemm_f16", 16384, 16384, 16384) == 256 and q("hgemm_f16", 4096, 4096, 4096) == 0
    assert L.cuxb_workspace_bytes(None, b"nope", 1, 1, 1, C.byref(b)) != 0


def test_library_is_sm100a_tensor_code():
    """The built library really is what DESIGN.md says: sm_100a cubins only, tcgen05 MMAs (UTCHMMA / UTCIMMA), TMEM
    loads (LDTM), TMA loads and stores (UTMALDG / UTMASTG), packed FFMA2 in the SGEMM -- read from its SASS."""
    import shutil
    import subprocess
    import sys
    if shutil.which("cuobjdump") is None and not os.path.exists("/usr/local/cuda/bin/cuobjdump"):
        pytest.skip("cuobjdump not available")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.run([sys.executable, os.path.join(root, "tools", "sass_opcodes.py")], check=True, capture_output=True, timeout=300)
    txt = open(os.path.join(root, "profiles", "sass_opcodes_r02.txt")).read()
    assert "cubin architectures: sm_100a\n" in txt
    for op in ("UTCHMMA.2CTA", "UTCIMMA.2CTA", "LDTM", "UTMALDG", "UTMASTG", "UBLKCP", "FFMA2"):
        assert op in txt, op
    total = txt.strip().splitlines()[-1]
    assert total.startswith("# whole library:") and "UTCHMMA" in total and "FFMA2 2624" in total


def test_bench_descriptions_are_stable_across_gpu_counts():
    """The driver lines the per-N bench values up by `metric`; both arms print the same config."""
    import importlib.util
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("bench", os.path.join(root, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    for name in bench.ALL:
        metrics = {bench.describe(name, n, "strong")["metric"] for n in (1, 2, 4, 8)}
        assert len(metrics) == 1, name
        assert "1/8 of it per GPU" in bench.describe(name, 8, "strong")["config"]["workload"]
        assert "per GPU" in bench.describe(name, 8, "weak")["config"]["workload"]
