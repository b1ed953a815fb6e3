"""Turns an `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv` launch list of one
training step into profiles/r02_dram_traffic.json (per kernel: launches, DRAM bytes per launch, time per launch).
    python tools/ncu_traffic.py <launches.csv> <bn_freeze 0|1> [out.json]"""
import collections
import csv
import json
import os
import sys

src, freeze = sys.argv[1], int(sys.argv[2])
out = sys.argv[3] if len(sys.argv) > 3 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                         "profiles", "r02_dram_traffic.json")
rows = list(csv.reader(open(src)))
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
h = rows[hdr]
ki, mi, ui, vi, idi = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Unit"), h.index("Metric Value"), h.index("ID")
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}
acc = collections.defaultdict(lambda: {"ids": set(), "bytes": 0.0, "us": 0.0})
for r in rows[hdr + 1:]:
    if len(r) <= vi:
        continue
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    name = r[ki].split("(")[0].replace("void ", "").replace("asf::", "").split("<")[0]
    a = acc[name]
    a["ids"].add(r[idi])
    if r[mi].startswith("dram__bytes"):
        a["bytes"] += v * scale.get(r[ui], 1.0)
    elif r[mi].startswith("gpu__time_duration"):
        a["us"] += v * scale.get(r[ui], 1.0)
d = {}
if os.path.exists(out):
    d = json.load(open(out))
d["bn_freeze_%d" % freeze] = {k: {"launches": len(a["ids"]), "dram_bytes_per_launch": a["bytes"] / max(1, len(a["ids"])),
                                   "us_per_launch": a["us"] / max(1, len(a["ids"]))}
                              for k, a in sorted(acc.items(), key=lambda x: -x[1]["us"]) if a["us"] > 0}
json.dump(d, open(out, "w"), indent=1)
for k, a in list(d["bn_freeze_%d" % freeze].items())[:12]:
    print("%-36s n=%4d  %10.0f B/launch  %8.1f us/launch" % (k, a["launches"], a["dram_bytes_per_launch"], a["us_per_launch"]))
