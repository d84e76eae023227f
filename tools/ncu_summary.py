"""Summarise an `ncu --csv` launch list: per kernel count, mean duration, DRAM read/write bytes."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
H, data = rows[hdr], rows[hdr + 1:]
ki, mi, vi, ii = H.index("Kernel Name"), H.index("Metric Name"), H.index("Metric Value"), H.index("ID")
per = collections.OrderedDict()
for r in data:
    if len(r) > vi:
        per.setdefault(r[ii], {"k": r[ki]})[r[mi]] = float(r[vi].replace(",", ""))
agg = collections.OrderedDict()
for d in per.values():
    if not d["k"].startswith(("ps::", "void ps::")) and len(sys.argv) < 3:
        continue
    a = agg.setdefault(d["k"][:90], [0, 0, 0, 0])
    a[0] += 1
    a[1] += d.get("gpu__time_duration.sum", 0)
    a[2] += d.get("dram__bytes_read.sum", 0)
    a[3] += d.get("dram__bytes_write.sum", 0)
print(f"{'n':>4} {'mean us':>9} {'rd MB':>8} {'wr MB':>8}  kernel")
for k, a in agg.items():
    print(f"{a[0]:4d} {a[1] / a[0] / 1e3:9.1f} {a[2] / a[0] / 1e6:8.1f} {a[3] / a[0] / 1e6:8.1f}  {k}")
