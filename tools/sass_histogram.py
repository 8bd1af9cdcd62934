#!/usr/bin/env python
"""Per-kernel SASS opcode histogram of libivx.so (cuobjdump -sass), for the opcodes that prove the
Blackwell-native path: UTCHMMA / UTCQMMA (tcgen05.mma), LDTM / STTM (tcgen05.ld / st), UTMALDG /
UTMASTG (TMA load / store), UTCBAR (tcgen05.commit), SYNCS (mbarrier), plus FFMA / HMMA for contrast.

    python tools/sass_histogram.py > profiles/sass_opcodes_r2.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "imvoxelnet_b200", "libivx.so")
WATCH = ["UTCHMMA", "UTCQMMA", "UTCIMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTCBAR", "SYNCS", "HMMA", "FFMA",
         "LDG", "STG", "RED", "ATOM", "MULTIMEM"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels, name = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(.*", "", name)
            kernels[name] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m and name:
            kernels[name][m.group(1)] += 1
            kernels[name]["_total"] += 1
    print("# %s, %d kernels; columns: %s" % (os.path.basename(LIB), len(kernels), " ".join(WATCH)))
    for k, c in kernels.items():
        cells = " ".join("%s=%d" % (w, sum(v for op, v in c.items() if op.startswith(w))) for w in WATCH
                         if any(op.startswith(w) for op in c))
        print("%-70s total=%-6d %s" % (k[:70], c["_total"], cells))


if __name__ == "__main__":
    sys.exit(main())
