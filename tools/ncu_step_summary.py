"""Per-launch table + per-kernel summary from tools/ncu_step_table.sh's raw-page CSV of one eager training step.
    python tools/ncu_step_summary.py <raw.csv> [top_n]"""
import collections
import csv
import sys

COLS = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "rd_MB"), ("dram__bytes_write.sum", "wr_MB"),
        ("sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tc%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%"),
        ("launch__registers_per_thread", "regs"), ("smsp__inst_executed.sum", "Minst")]
UNIT = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "byte": 1e-6, "Kbyte": 1e-3,
        "Mbyte": 1.0, "Gbyte": 1e3}
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 12]
h, units, body = rows[0], rows[1], rows[2:]
ix = {n: h.index(n) for n, _ in COLS if n in h}
ki, gi = h.index("Kernel Name"), h.index("Grid Size")


def val(r, n):
    v = float(r[ix[n]].replace(",", "") or 0)
    if n == "smsp__inst_executed.sum":
        return v / 1e6
    return v * UNIT.get(units[ix[n]], 1.0)


recs = []
for r in body:
    name = r[ki].split("(")[0].replace("void ", "").replace("asf::", "")[:40]
    recs.append((name, r[gi], [val(r, n) for n, _ in COLS if n in ix]))
names = [s for n, s in COLS if n in ix]
tot = sum(v[0] for _, _, v in recs)
print("one eager training step at the bench shape: %d launches, %.1f us summed (serialised, cold caches)" % (len(recs), tot))
agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0, 0.0])
for name, _, v in recs:
    a = agg[name]
    a[0] += 1
    a[1] += v[0]
    a[2] += v[1] + v[2]
    a[3] += v[3] * v[0]
    a[4] += v[6] * v[0]
print("\n%-42s %5s %10s %6s %10s %12s %8s" % ("kernel", "n", "us", "share", "dram_MB", "tc% (t-wtd)", "dram%"))
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%-42s %5d %10.1f %5.1f%% %10.1f %12.1f %8.1f" % (k, a[0], a[1], 100 * a[1] / tot, a[2], a[3] / a[1], a[4] / a[1]))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
print("\n%d longest launches (launch order index first)" % top)
print("%4s %-40s %-14s " % ("#", "kernel", "grid") + " ".join("%8s" % s for s in names))
for i, (name, grid, v) in sorted(enumerate(recs), key=lambda x: -x[1][2][0])[:top]:
    print("%4d %-40s %-14s " % (i, name, grid) + " ".join("%8.1f" % x for x in v))
