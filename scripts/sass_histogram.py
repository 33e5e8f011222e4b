#!/usr/bin/env python
"""Per-kernel SASS opcode histogram of nerf_b200/libnerf_b200.so (cuobjdump -sass): which kernels carry the Blackwell
instructions (UTCHMMA = tcgen05.mma, UTCBAR = tcgen05.commit, LDTM/STTM = tcgen05.ld/st, UBLKCP/UBLKPF = bulk TMA copy /
prefetch, UTMALDG = tensor-map TMA, SYNCS = mbarrier, FFMA2/FADD2 = packed fp32) and what the rest is made of.

    python scripts/sass_histogram.py > profiles/r02_sass_opcodes.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "nerf_b200", "libnerf_b200.so")
KEY = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTCATOMSWS", "UBLKCP", "UBLKPF", "UBLKRED", "UTMALDG", "UTMASTG", "SYNCS", "REDG", "RED",
       "ATOMG", "MUFU", "FFMA2", "FADD2", "FMUL2", "HMNMX2", "F2FP", "FFMA", "IMAD", "LDS", "STS", "LDG", "STG", "SHFL", "BAR", "ELECT", "NANOSLEEP"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = kernels.setdefault(m.group(1), collections.Counter()); continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m and cur is not None:
            cur[m.group(1)] += 1
    demangle = subprocess.run(["cu++filt"] + list(kernels), capture_output=True, text=True).stdout.splitlines() if kernels else []
    names = dict(zip(kernels, demangle)) if len(demangle) == len(kernels) else {k: k for k in kernels}
    total = collections.Counter()
    print(f"# cuobjdump -sass {os.path.relpath(SO, ROOT)}: opcode counts per kernel (static instructions)")
    print(f"# {'kernel':70s} {'instr':>7s}  " + " ".join(f"{k}" for k in KEY if any(c[k] for c in kernels.values())))
    cols = [k for k in KEY if any(c[k] for c in kernels.values())]
    for mangled, c in sorted(kernels.items(), key=lambda kv: -sum(kv[1].values())):
        n = names[mangled].rsplit("(", 1)[0].replace("(bool)", "").replace("void ", "")[:70]
        print(f"{n:72s} {sum(c.values()):7d}  " + " ".join(f"{c[k]:>{len(k)}d}" for k in cols))
        total.update(c)
    print(f"{'TOTAL':72s} {sum(total.values()):7d}  " + " ".join(f"{total[k]:>{len(k)}d}" for k in cols))
    print("# tcgen05 / TMA evidence (whole library): " + ", ".join(f"{k}={total[k]}" for k in ("UTCHMMA", "UTCBAR", "LDTM", "UBLKCP", "UBLKPF", "UTMALDG", "SYNCS")))


if __name__ == "__main__":
    sys.exit(main())
