"""Summarise an ncu --set full report (.ncu-rep) as one CSV line per profiled launch.
Usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x_summary.csv"""
import csv, subprocess, sys, io

COLS = [
    ("ms", "gpu__time_duration.sum", 1e-6),
    ("grid", "launch__grid_size", 1),
    ("regs", "launch__registers_per_thread", 1),
    ("smem_dyn_KB", "launch__shared_mem_per_block_dynamic", 1 / 1024),
    ("tensor_pct", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", 1),
    ("tensor_rt_pct", "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", 1),
    ("sm_busy_pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed", 1),
    ("dram_pct", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1),
    ("dram_rd_MB", "dram__bytes_read.sum", 1e-6),
    ("dram_wr_MB", "dram__bytes_write.sum", 1e-6),
    ("l2_pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed", 1),
    ("l1_pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", 1),
    ("l2_rd_MB_from_tex", "lts__t_bytes_srcunit_tex_op_read.sum", 1e-6),
    ("smem_wavefronts_M", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", 1e-6),
    ("sm_mhz", "sm__cycles_elapsed.avg.per_second", 1e-6),
    ("warps_active_pct", "sm__warps_active.avg.pct_of_peak_sustained_active", 1),
]


def main(path, extra=()):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    w = csv.writer(sys.stdout)
    cols = [c for c in COLS if c[1] in idx]
    w.writerow(["id", "kernel"] + [c[0] for c in cols])
    for r in data:
        out = [r[idx["ID"]], r[idx["Kernel Name"]].split("(")[0][:60]]
        for name, key, scale in cols:
            v = r[idx[key]].replace(",", "")
            u = units[idx[key]]
            try:
                f = float(v)
                if name == "ms":
                    f *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1e-6)
                elif "MB" in name:
                    f *= {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1e-6)
                elif name == "smem_dyn_KB":
                    f *= {"byte": 1 / 1024, "Kbyte": 1.0}.get(u, 1 / 1024)
                elif name == "sm_mhz":
                    f *= {"hz": 1e-6, "Khz": 1e-3, "Mhz": 1.0, "Ghz": 1e3}.get(u, 1e-6)
                else:
                    f *= scale
                out.append(f"{f:.4g}")
            except ValueError:
                out.append(v)
        w.writerow(out)


if __name__ == "__main__":
    main(sys.argv[1])
