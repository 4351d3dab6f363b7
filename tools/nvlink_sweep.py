"""NVLink pair bandwidth by transfer style and bytes in flight (cuxb_probe_nvlink), GPU 1 -> GPU 0.
    python tools/nvlink_sweep.py            (needs 2 visible GPUs, single process)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hpc_b200 import capi
ctx = capi.Context(0)
for mode in ("pull", "push", "memcpy"):
    print("%-8s %7.1f GB/s" % (mode, ctx.probe_nvlink(1, mode, 0, 512 << 20, 3)), flush=True)
for depth in (1, 2, 3, 4, 6, 8, 12):
    print("bulk depth %2d (%3d KiB in flight per SM) %7.1f GB/s" % (depth, depth * 16, ctx.probe_nvlink(1, "bulk", depth, 512 << 20, 3)), flush=True)
