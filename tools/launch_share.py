"""Aggregate an ncu launch list (--metrics gpu__time_duration.sum --csv) into per-kernel time shares of one
decoder pass, and (optionally) turn an ncu --set full summary (tools/ncu_summary.py) into the per-launch DRAM
traffic record bench.py reports as roofline.traffic.
    python tools/launch_share.py gpurun_out/x_launches.csv > profiles/x_launch_share.csv
    python tools/launch_share.py --traffic profiles/x_summary.csv 'conv_mma_kernel<128' CHUNKS [FIRST_ID LAST_ID] > profiles/roofline_traffic.json
(the id range picks the launches of one layer group, e.g. 21 38 = the 18 ResBlock convs of the 128-channel stage)"""
import csv, json, re, sys
from collections import OrderedDict


def short(name):
    name = name.replace("rvcdec::", "").replace("void ", "")
    return re.sub(r"\(.*$", "", name)


def shares(path):
    rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 14 and r[0].isdigit()]
    agg = OrderedDict()
    for r in rows:
        k = short(r[4])
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += float(r[14]) * 1e-6
    tot = sum(v[1] for v in agg.values())
    w = csv.writer(sys.stdout)
    w.writerow(["kernel", "launches", "total_ms", "share"])
    for k, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        w.writerow([k, n, f"{ms:.3f}", f"{ms / tot:.4f}"])
    w.writerow(["TOTAL", len(rows), f"{tot:.3f}", "1.0"])


def traffic(path, pat, chunks, first=None, last=None):
    rows = list(csv.DictReader(open(path)))
    sel = [r for r in rows if pat in r["kernel"] and (first is None or first <= int(r["id"]) <= last)]
    b = [(float(r["dram_rd_MB"]) + float(r["dram_wr_MB"])) * 1e6 for r in sel]
    print(json.dumps({"kernel": pat, "launches": len(sel), "chunks": int(chunks), "frames_per_chunk": 3000,
                      "dram_bytes_per_launch": sum(b) / len(b), "ms_per_launch_under_ncu": sum(float(r["ms"]) for r in sel) / len(sel),
                      "source": path, "how": "mean of dram__bytes_read.sum + dram__bytes_write.sum over the launches of one decoder pass, ncu --set full --clock-control none"}))


if __name__ == "__main__":
    if sys.argv[1] == "--traffic":
        traffic(sys.argv[2], sys.argv[3], sys.argv[4], *(int(v) for v in sys.argv[5:7]))
    else:
        shares(sys.argv[1])
