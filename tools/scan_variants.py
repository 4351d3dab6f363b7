#!/usr/bin/env python3
"""Build libcuxb variants with different scan pipeline depths (tuning experiment).
usage: python tools/scan_variants.py   -> tools/_variants/libcuxb_<pf>_<gl>_<sl>.so"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hpc_b200 import build as b
OUT = os.path.join(ROOT, "tools", "_variants")
os.makedirs(OUT, exist_ok=True)
for pf, gl, sl in [(2, 2, 3), (3, 1, 2), (3, 2, 3), (4, 1, 2), (2, 1, 2)]:
    so = os.path.join(OUT, "libcuxb_%d_%d_%d.so" % (pf, gl, sl))
    b.build(verbose=False, extra_flags=["-DCUXB_SCAN_PREFETCH=%d" % pf, "-DCUXB_SCAN_GATHER_LAG=%d" % gl,
                                        "-DCUXB_SCAN_STORE_LAG=%d" % sl], out=so, build_dir=so + ".build")
    print(so)
