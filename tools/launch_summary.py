"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
h = rows[hdr]
ki, vi = h.index("Kernel Name"), h.index("Metric Value")
mi, ui = h.index("Metric Name"), h.index("Metric Unit")
SCALE = {"ns": 1.0, "nsecond": 1.0, "us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6}
d = collections.defaultdict(lambda: [0, 0.0])
for r in rows[hdr + 1:]:
    if len(r) <= vi or not r[mi].startswith("gpu__time_duration"):
        continue
    try:
        v = float(r[vi].replace(",", "")) * SCALE.get(r[ui], 1.0)
    except ValueError:
        continue
    name = r[ki].split("(")[0].replace("void ", "")[:70]
    d[name][0] += 1
    d[name][1] += v
tot = sum(v[1] for v in d.values())
print("%-72s %6s %10s %6s" % ("kernel", "count", "total_us", "share"))
for k, v in sorted(d.items(), key=lambda x: -x[1][1]):
    print("%-72s %6d %10.1f %5.1f%%" % (k, v[0], v[1] / 1e3, 100 * v[1] / tot))
print("%-72s %6d %10.1f" % ("TOTAL", sum(v[0] for v in d.values()), tot / 1e3))
