#!/usr/bin/env python3
"""Build recipe for the parity oracle (TEST INFRASTRUCTURE, not product code).

  python oracle/build.py            # builds both targets that can be built
  python oracle/build.py oracle     # oracle/liboracle.so   (CPU restatement, plain C)
  python oracle/build.py ref        # oracle/_ref/libcuxref.so (the reference's own host code)

`ref` needs /root/reference (present in the authoring container only).  It
compiles the reference's standalone samples from where they lie and extracts
the cux host kernels by line range into oracle/_ref/*.inc (git-ignored; the
cux translation units themselves only build with MSVC).  Nothing
from the reference is written anywhere else.  The resulting .so travels to the
GPU box with the snapshot (oracle/_ref is git-ignored but NOT gpurun-ignored).

Flags: -ffp-contract=off everywhere, so scalar `t += a*b` rounds twice exactly
as written (cux targets MSVC /fp:precise, which never contracts); explicit
_mm256_fmadd_ps intrinsics in the reference keep their single rounding.
ISA: the oracle uses baseline x86-64 only; the reference objects are compiled
with the ISA flags their intrinsics need and the shim checks CPUID at run time
before entering an AVX-512 variant.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("CUXB_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")
CUDA_INC = os.environ.get("CUDA_HOME", "/usr/local/cuda") + "/include"


def run(cmd):
    print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources if os.path.exists(s))


def build_oracle(force=False):
    src = os.path.join(HERE, "oracle.c")
    so = os.path.join(HERE, "liboracle.so")
    if force or newer(so, [src, __file__]):
        run(["gcc", "-std=c11", "-O2", "-fopenmp", "-ffp-contract=off", "-fno-fast-math",
             "-fvisibility=hidden", "-fPIC", "-shared", "-o", so, src, "-lm"])
    return so


def _extract(rel, first, last, dst):
    """Copy lines [first,last] of a reference file into oracle/_ref (build intermediate)."""
    with open(os.path.join(REF, rel), "r", errors="replace") as f:
        lines = f.readlines()
    with open(os.path.join(OUT, dst), "w") as f:
        f.write("// build intermediate extracted from %s:%d-%d -- do not commit\n" % (rel, first, last))
        f.writelines(lines[first - 1:last])


def build_ref(force=False):
    """oracle/_ref/libcuxref.so: the PARITY build (-O2 -ffp-contract=off: scalar code rounds as written).
    oracle/_ref/libcuxref_fast.so: the same sources at -O3 for an x86-64-v3 host (AVX2 + FMA; contraction
    allowed) -- only ever TIMED (bench.py's cpu_baseline / --impl reference legs), never compared against.
    -march=native is deliberately not used: the library is built in this container and run on the GPU box,
    whose CPU need not be the same model; the AVX-512 variants keep the per-file ISA flags their intrinsics
    need and are entered only after a CPUID check."""
    if not os.path.isdir(REF):
        print("reference tree %s not present: keeping prebuilt oracle/_ref as is" % REF)
        return None
    os.makedirs(OUT, exist_ok=True)
    shim = os.path.join(HERE, "ref_shim.cpp")
    out = None
    for name, opt in (("libcuxref.so", ["-O2", "-ffp-contract=off"]),
                      ("libcuxref_fast.so", ["-O3", "-march=x86-64-v3", "-funroll-loops"])):
        so = os.path.join(OUT, name)
        out = out or so
        if not (force or newer(so, [shim, __file__])):
            continue
        _extract("0-frameworks/cux/src/operator/gemm.cpp", 11, 275, "gen_cux_gemm.inc")
        _extract("0-frameworks/cux/src/operator/dot_product.cpp", 5, 13, "gen_cux_dot.inc")
        _extract("0-frameworks/cux/src/operator/nrm2.cpp", 5, 16, "gen_cux_nrm2.inc")
        _extract("cuda/gemm_fp16_wmma.cu", 75, 90, "gen_hgemm_host.inc")
        _extract("cuda/reduce_fp32.cu", 26, 31, "gen_reduce_host.inc")
        common = opt + ["-fopenmp", "-fPIC", "-w", "-fpermissive"]
        objs = []
        # Standalone samples, compiled from where they lie; only main() is renamed.
        for rel, main, extra in [
            ("simd/x86/linear.cpp", "ref_linear_main",
             ["-std=c++17", "-mavx2", "-mfma", "-mf16c", "-mavx512f", "-mavx512bw", "-mavx512vl",
              "-mavx512vnni", "-mavx512bf16", "-I" + REF]),
            ("simd/x86/vector_scan.cpp", "ref_scan_main", ["-std=c++17", "-mavx2"]),
            ("openmp/alg_matrix_multiply.cpp", "ref_ompmm_main", ["-std=c++14"]),
            # the sample relies on MSVC's headers for memset / clock: supply them on the command line
            ("std/alg_vector_dot_product.cpp", "ref_stddot_main",
             ["-std=c++14", "-pthread", "-include", "cstring", "-include", "ctime"]),
        ]:
            obj = os.path.join(OUT, os.path.basename(rel) + ".o")
            run(["g++"] + common + extra + ["-Dmain=" + main, "-c", os.path.join(REF, rel), "-o", obj])
            objs.append(obj)
        obj = os.path.join(OUT, "ref_shim.o")
        run(["g++", "-std=c++14", "-mavx2", "-mfma"] + common +
            ["-I" + OUT, "-I" + CUDA_INC, "-c", shim, "-o", obj])
        objs.append(obj)
        run(["g++", "-shared", "-fopenmp", "-pthread", "-o", so] + objs)
        # keep only the library: intermediates (objects, extracted line ranges) are removed
        for f in os.listdir(OUT):
            if f.endswith(".o") or f.endswith(".inc"):
                os.remove(os.path.join(OUT, f))
    return out


def main(argv):
    what = argv[1] if len(argv) > 1 else "all"
    if what in ("oracle", "all"):
        build_oracle(force="--force" in argv)
    if what in ("ref", "all"):
        build_ref(force="--force" in argv)


if __name__ == "__main__":
    main(sys.argv)
