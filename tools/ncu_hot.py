"""Top SASS instructions by warp-stall samples for one profiled launch.
Usage: python tools/ncu_hot.py report.ncu-rep <launch-index-within-report (1-based)> [topN]"""
import csv, io, subprocess, sys
path, k = sys.argv[1], int(sys.argv[2]); top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
raw = subprocess.run(["ncu", "-i", path, "--page", "source", "--print-source", "sass", "--csv", "--launch-skip", str(k - 1), "--launch-count", "1"], capture_output=True, text=True).stdout
lines = raw.splitlines()
print(lines[0][:160])
rows = list(csv.reader(io.StringIO("\n".join(lines[1:]))))
hdr = rows[0]; idx = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = [r for r in rows[1:] if len(r) == len(hdr) and r[idx["# Samples"]].strip().isdigit()]
tot = sum(int(r[idx["# Samples"]]) for r in data)
print("total samples", tot, "instructions", len(data))
order = sorted(range(len(data)), key=lambda i: -int(data[i][idx["# Samples"]]))[:top]
for i in sorted(order):
    r = data[i]; s = int(r[idx["# Samples"]])
    st = sorted(((int(r[idx[h]]), h[6:]) for h in stalls), reverse=True)[:2]
    print(f"{i:5d} {100*s/tot:5.1f}% exec={r[idx['Instructions Executed']]:>9} {r[idx['Source']].strip()[:90]:<90} {st}")
