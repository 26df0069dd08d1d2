#!/usr/bin/env python
"""One line per captured launch of an ncu report: duration, DRAM traffic/throughput, issue utilisation, occupancy.
usage: tools/ncu_summary.py report.ncu-rep"""
import csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
cols = {"t_us": "gpu__time_duration.sum", "rd_MB": "dram__bytes_read.sum", "wr_MB": "dram__bytes_write.sum",
        "dram%": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "issue%": "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "occ%": "sm__warps_active.avg.pct_of_peak_sustained_active", "regs": "launch__registers_per_thread",
        "l2rd_MB": "lts__t_sectors_srcunit_tex_op_read.sum", "fma%": "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"}
def conv(v, u, name):
    v = float(v.replace(",", ""))
    if name == "t_us": return v * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(u.replace("second", "s").replace("nsecond","ns"), 1)
    if name in ("rd_MB", "wr_MB"): return v * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1, "Gbyte": 1e3}.get(u, 1)
    if name == "l2rd_MB": return v * 32e-6
    return v
print(f"{'kernel':34s}" + "".join(f"{k:>9s}" for k in cols))
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")].split("(")[0][:34]
    vals = []
    for k, m in cols.items():
        if m in hdr:
            i = hdr.index(m); vals.append(conv(r[i], units[i], k))
        else: vals.append(float("nan"))
    print(f"{name:34s}" + "".join(f"{v:9.1f}" for v in vals))
