"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) per kernel: count, total ms, mean us."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
hdr = rows[hi]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
gi = hdr.index("Grid Size") if "Grid Size" in hdr else None
agg = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) <= vi:
        continue
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    key = r[ki][:70] + (" grid=" + r[gi] if gi is not None else "")
    a = agg.setdefault(key, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
print(f"{'total ms':>10} {'n':>5} {'mean us':>10} {'share':>6}  kernel")
for k, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{t / 1e6:10.3f} {c:5d} {t / c / 1e3:10.1f} {t / tot * 100:5.1f}%  {k}")
