#!/usr/bin/env python3
"""Build the cux-compatible C++ host layer with g++ (no nvcc, no CUDA headers):
    hpc_b200/cux/libcux_host.so   the framework (Executor, ops, Array4D, ...)
    hpc_b200/cux/cux_demo         example/demo.cpp + example/user_plugins.cpp
    hpc_b200/cux/cux_host_test    tests/host_test.cpp (runs without a GPU)
    hpc_b200/cux/cux_device_test  tests/device_test.cpp (the reference's DataTest cases; needs a B200)
All link against hpc_b200/libcuxb.so through include/cuxb.h only.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
ROOT = os.path.dirname(PKG)
INC = ["-I" + os.path.join(HERE, "include"), "-I" + os.path.join(ROOT, "include")]
CXX = ["g++", "-std=c++17", "-O2", "-fPIC", "-Wall", "-Wno-unused-function"]
LINK = ["-L" + PKG, "-lcuxb", "-Wl,-rpath," + PKG, "-Wl,-rpath," + HERE]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _all_headers():
    out = [os.path.join(ROOT, "include", "cuxb.h"), os.path.abspath(__file__), os.path.join(PKG, "libcuxb.so")]
    for dp, _, fs in os.walk(os.path.join(HERE, "include")):
        out += [os.path.join(dp, f) for f in fs]
    return out


def run(cmd):
    print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def build(force=False):
    hdrs = _all_headers()
    srcs = [os.path.join(HERE, "src", f) for f in sorted(os.listdir(os.path.join(HERE, "src"))) if f.endswith(".cpp")]
    lib = os.path.join(HERE, "libcux_host.so")
    if force or _stale(lib, srcs + hdrs):
        run(CXX + INC + ["-shared", "-o", lib] + srcs + LINK)
    demo = os.path.join(HERE, "cux_demo")
    dsrc = [os.path.join(HERE, "example", "demo.cpp"), os.path.join(HERE, "example", "user_plugins.cpp")]
    if force or _stale(demo, dsrc + hdrs + [lib]):
        run(CXX + INC + ["-o", demo] + dsrc + ["-L" + HERE, "-lcux_host"] + LINK)
    test = os.path.join(HERE, "cux_host_test")
    tsrc = [os.path.join(HERE, "tests", "host_test.cpp")]
    if force or _stale(test, tsrc + hdrs + [lib]):
        run(CXX + INC + ["-o", test] + tsrc + ["-L" + HERE, "-lcux_host"] + LINK)
    dtest = os.path.join(HERE, "cux_device_test")
    dtsrc = [os.path.join(HERE, "tests", "device_test.cpp")]
    if force or _stale(dtest, dtsrc + hdrs + [lib]):
        run(CXX + INC + ["-o", dtest] + dtsrc + ["-L" + HERE, "-lcux_host"] + LINK)
    return lib, demo, test


if __name__ == "__main__":
    build(force="--force" in sys.argv)
