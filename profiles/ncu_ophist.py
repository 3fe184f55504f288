"""Opcode histogram (weighted by executed warp instructions) from `ncu --page source --csv` output."""
import csv
import sys
from collections import Counter

rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ia, ie, isamp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
data = []
for r in rows[hi + 1:]:
    if len(r) > ie and r[ie].isdigit():
        data.append((r[ia], int(r[ie]), int(r[isamp])))
tot = sum(d[1] for d in data)
print("total warp instructions", tot, "distinct SASS", len(data))
c, s = Counter(), Counter()
for src, e, sm in data:
    t = src.split()
    op = t[1] if t[0].startswith("@") else t[0]
    op = op.split(".")[0]
    c[op] += e
    s[op] += sm
ts = sum(s.values())
for op, e in c.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 25):
    print(f"{op:10s} {e / tot * 100:5.1f}% of instructions   {s[op] / max(ts, 1) * 100:5.1f}% of stall samples")
