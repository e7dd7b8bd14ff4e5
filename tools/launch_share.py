"""Per-kernel totals of an ncu --metrics gpu__time_duration.sum launch list (second half of the list = the timed call when the
script ran the same work twice): python tools/launch_share.py list.csv [all]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1], errors="ignore")))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
H = rows[hdr]
ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
data = [r for r in rows[hdr + 1:] if len(r) > vi and r[vi].replace(",", "").replace(".", "").isdigit()]
if len(sys.argv) < 3:
    data = data[len(data) // 2:]
agg = collections.OrderedDict()
for r in data:
    v = float(r[vi].replace(",", ""))
    v = v / 1e3 if r[ui] in ("ns", "nsecond") else v
    a = agg.setdefault(r[ki][:48], [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
for k, a in agg.items():
    print(f"{k:50s} n={a[0]:4d} {a[1] / 1e3:9.3f} ms {100 * a[1] / tot:5.1f} %")
print(f"total {tot / 1e3:.3f} ms over {len(data)} launches")
