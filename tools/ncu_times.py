"""Sums gpu__time_duration per kernel name from an `ncu --csv --metrics gpu__time_duration.sum` log."""
import csv
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith("=="))]
hdr = rows[0]
ki, mi, vi = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
acc = OrderedDict()
for r in rows[1:]:
    if r[mi] != "gpu__time_duration.sum":
        continue
    name = r[ki].split("(")[0].replace("<unnamed>::", "")
    t, n = acc.get(name, (0.0, 0))
    acc[name] = (t + float(r[vi].replace(",", "")), n + 1)
tot = sum(t for t, _ in acc.values())
for k, (t, n) in acc.items():
    print(f"{k:24s} launches {n:3d}  total {t / 1e6:9.3f} ms  mean {t / n / 1e3:10.1f} us  share {t / tot:6.1%}")
