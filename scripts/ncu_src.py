#!/usr/bin/env python
"""Top stall sites of one kernel from `ncu --page source --csv` output.
   ncu -i rep --page source --csv --kernel-name regex:X > f.csv ; python scripts/ncu_src.py f.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
h = rows[hi]
si, ss, ie = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
data = []
for r in rows[hi + 1:]:
    try:
        data.append((int(r[ss] or 0), int(r[ie] or 0), r[si], r))
    except (ValueError, IndexError):
        pass
tot = sum(d[0] for d in data)
print("total samples", tot, "total warp instr", sum(d[1] for d in data))
stall_cols = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
for smp, ex, src, r in sorted(data, key=lambda d: -d[0])[:n]:
    st = sorted(((int(r[i] or 0), h[i][6:]) for i in stall_cols), reverse=True)[:2]
    print(f"{100*smp/max(tot,1):5.1f}% ex={ex:9d} {src[:70]:70s} {st}")
