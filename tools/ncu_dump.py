"""Dump SASS with samples for one launch. Usage: ncu_dump.py report idx(1-based) > file"""
import csv, io, subprocess, sys
path, k = sys.argv[1], int(sys.argv[2])
raw = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--launch-skip", str(k - 1), "--launch-count", "1"], capture_output=True, text=True).stdout
lines = raw.splitlines()
rows = list(csv.reader(io.StringIO("\n".join(lines[1:]))))
hdr = rows[0]; idx = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
for i, r in enumerate(rows[1:]):
    s = int(r[idx["# Samples"]])
    st = sorted(((int(r[idx[h]]), h[6:]) for h in stalls), reverse=True)[:2]
    st = [x for x in st if x[0] > 0]
    print(f"{i:5d} {s:6d} exec={r[idx['Instructions Executed']]:>9} {r[idx['Source']].strip()[:100]:<100} {st}")
