#!/usr/bin/env python
"""Per-kernel instruction census of libsdb200.so (cuobjdump -sass): how many tcgen05 MMA (UTCHMMA), TMEM load / store
(LDTM / STTM), TMA load / store (UTMALDG / UTMASTG, UBLKCP = 1-D bulk copy), legacy tensor-core (HMMA) and MUFU
instructions every kernel carries.  Evidence that the hot kernels are tcgen05 / TMA code and which ones still run on
mma.sync.

    python tools/sass_census.py > profiles/r02_sass_census.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "stable_diffusion_pytorch_b200", "libsdb200.so")
OPS = ["UTCHMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "HMMA", "MUFU", "SYNCS", "UCGABAR"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    counts, order, cur = collections.defaultdict(collections.Counter), [], None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            order.append(cur)
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m:
            counts[cur]["total"] += 1
            op = m.group(1)
            if op in OPS:
                counts[cur][op] += 1
    names = subprocess.run(["c++filt"] + order, capture_output=True, text=True).stdout.splitlines()
    print(f"# SASS census of {os.path.relpath(LIB, ROOT)} (sm_100a), {len(order)} kernels; columns = instruction counts")
    print(f"{'total':>7} " + " ".join(f"{o:>8}" for o in OPS) + "  kernel")
    rows = sorted(zip(order, names), key=lambda kv: (-counts[kv[0]]["UTCHMMA"], -counts[kv[0]]["HMMA"], kv[1]))
    for mangled, name in rows:
        c = counts[mangled]
        short = re.sub(r"\(.*", "", name).replace("void ", "")
        print(f"{c['total']:7d} " + " ".join(f"{c[o]:8d}" for o in OPS) + f"  {short[:110]}")
    tc = [n for m, n in rows if counts[m]["UTCHMMA"]]
    legacy = [n for m, n in rows if counts[m]["HMMA"] and not counts[m]["UTCHMMA"]]
    print(f"\n# {len(tc)} kernels issue tcgen05.mma (UTCHMMA); {len(legacy)} run on the legacy mma.sync pipe (HMMA)")


if __name__ == "__main__":
    sys.exit(main())
