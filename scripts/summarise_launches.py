"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel (count, total time, share)."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1], errors="replace")))
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
h, body = rows[hi], rows[hi + 1:]
ki, mi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
unit = None
for r in body:
    if len(r) <= mi:
        continue
    name = r[ki].split("(")[0][:80]
    agg[name][0] += 1
    agg[name][1] += float(r[mi].replace(",", ""))
    unit = r[ui]
scale = {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "nsecond": 1e-6, "ms": 1.0, "msecond": 1.0}.get(unit, 1e-6)
tot = sum(v[1] for v in agg.values())
print(f"# launches={sum(v[0] for v in agg.values())} total={tot * scale:.3f} ms (unit in file: {unit})")
print(f"{'count':>6} {'total ms':>12} {'share':>8}  kernel")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{v[0]:6d} {v[1] * scale:12.3f} {100 * v[1] / tot:7.2f}%  {k}")
