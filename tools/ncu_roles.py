"""Warp-stall samples of one profiled launch of a warp-specialised kernel, split by warp role.
The role regions are found from marker opcodes in the SASS (UBLKCP = TMA producer, UTCHMMA = MMA issuer,
LDTM / STTM = epilogue); every instruction between two markers of different roles is attributed to the role of the
nearest marker before it.  Usage: python tools/ncu_roles.py report.ncu-rep <launch (1-based)>"""
import csv, io, subprocess, sys
from collections import Counter, defaultdict

path, k = sys.argv[1], int(sys.argv[2])
raw = subprocess.run(["ncu", "-i", path, "--page", "source", "--print-source", "sass", "--csv", "--launch-skip", str(k - 1),
                      "--launch-count", "1"], capture_output=True, text=True).stdout
lines = raw.splitlines()
print(lines[0][:150])
rows = list(csv.reader(io.StringIO("\n".join(lines[1:]))))
hdr = rows[0]; idx = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[1:] if len(r) == len(hdr) and r[idx["# Samples"]].strip().isdigit()]
data = data[: len(data) // 2] if len(data) > 1 and data[0][idx["Address"]] == data[len(data) // 2][idx["Address"]] else data
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]


def marker(src):
    if "UBLKCP" in src: return "producer"
    if "UTCHMMA" in src or "UTCBAR" in src: return "mma"
    if "LDTM" in src or "STTM" in src: return "epilogue"
    return None


role, roles = "prologue", []
for r in data:
    mk = marker(r[idx["Source"]])
    if mk: role = mk
    roles.append(role)
# an instruction before the first marker of a region belongs to that region too: smooth backwards over short gaps
tot = sum(int(r[idx["# Samples"]]) for r in data)
by_role = defaultdict(Counter)
ops = defaultdict(Counter)
for r, ro in zip(data, roles):
    s = int(r[idx["# Samples"]])
    by_role[ro]["samples"] += s
    for h in stalls:
        by_role[ro][h[6:]] += int(r[idx[h]])
    op = r[idx["Source"]].split()
    op = next((t for t in op if not t.startswith("@")), "?").split(".")[0]
    ops[ro][op] += s
print("total samples", tot)
for ro, c in by_role.items():
    s = c.pop("samples")
    top = ", ".join(f"{n} {100 * v / max(s, 1):.0f}%" for n, v in c.most_common(6))
    print(f"{ro:9s} {100 * s / tot:5.1f}% of samples | {top}")
    print("          hottest opcodes:", ", ".join(f"{n} {100 * v / max(s, 1):.0f}%" for n, v in ops[ro].most_common(8)))
