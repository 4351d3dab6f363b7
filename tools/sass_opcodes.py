#!/usr/bin/env python3
"""Per-kernel SASS opcode histogram of hpc_b200/libcuxb.so -> profiles/sass_opcodes_r02.txt.

    python tools/sass_opcodes.py

Evidence that the kernels are what DESIGN.md says they are: UTCHMMA / UTCIMMA (tcgen05.mma kind::f16 / i8),
LDTM (tcgen05.ld), UTMALDG / UTMASTG (TMA tensor loads / stores), UBLKCP (bulk copies), UTCBAR (tcgen05.commit),
SYNCS (mbarrier), FFMA2 (packed fp32 fma), LDGSTS (cp.async), REDUX / SHFL.  Run by __graft_entry__.build();
needs only cuobjdump (no GPU)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "hpc_b200", "libcuxb.so")
OUT = os.path.join(ROOT, "profiles", "sass_opcodes_r02.txt")
WATCH = ["UTCHMMA", "UTCIMMA", "UTCQMMA", "LDTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "SYNCS", "FFMA2", "FFMA", "HMMA",
         "IMMA", "LDGSTS", "LDG", "STG", "LDS", "STS", "SHFL", "ATOM", "RED", "BAR", "MEMBAR", "ERRBAR", "CCTL"]


def main():
    cuobjdump = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "bin", "cuobjdump")
    txt = subprocess.run([cuobjdump, "-sass", SO], capture_output=True, text=True, check=True).stdout
    arch = sorted(set(re.findall(r"arch = (sm_\w+)", txt)))
    kernels, cur = collections.OrderedDict(), None
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)((?:\.[A-Z0-9_]+)*)", line)
        if m and cur is not None:
            kernels[cur][m.group(1)] += 1
            if m.group(1) in ("UTCHMMA", "UTCIMMA", "UTMALDG", "UTMASTG", "UTCBAR", "LDTM", "UBLKCP") and m.group(2):
                kernels[cur][m.group(1) + m.group(2)] += 1
    demangle = subprocess.run(["c++filt"], input="\n".join(kernels), capture_output=True, text=True).stdout.splitlines()
    with open(OUT, "w") as f:
        f.write("# SASS opcode counts per kernel of hpc_b200/libcuxb.so (cuobjdump -sass); cubin architectures: %s\n" % ", ".join(arch))
        f.write("# produced by tools/sass_opcodes.py; static instruction counts, not execution counts\n\n")
        total = collections.Counter()
        for (name, cnt), nice in zip(kernels.items(), demangle):
            short = re.sub(r"\(.*", "", nice)
            f.write("%s\n" % short)
            keys = [k for k in cnt if k.split(".")[0] in WATCH]
            keys.sort(key=lambda k: (WATCH.index(k.split(".")[0]), k))
            f.write("    total %d instructions; " % sum(v for k, v in cnt.items() if "." not in k))
            f.write(", ".join("%s %d" % (k, cnt[k]) for k in keys) + "\n")
            for k in keys:
                if "." not in k:
                    total[k] += cnt[k]
        f.write("\n# whole library: " + ", ".join("%s %d" % (k, total[k]) for k in WATCH if total[k]) + "\n")
    print(OUT)


if __name__ == "__main__":
    sys.exit(main())
