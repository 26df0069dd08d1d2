#!/usr/bin/env python
"""Build profiles/ncu_traffic.json (what bench.py quotes as `roofline.traffic` / `ncu_pipe_utilisation`) from one
`ncu --set full` report of a single step of the bench workload.
usage: tools/ncu_traffic.py report.ncu-rep capture_name > profiles/ncu_traffic.json"""
import csv, io, json, subprocess, sys

rep, capture = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]


def col(r, name):
    i = hdr.index(name)
    v = float(r[i].replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(units[i], 1)


def stage_of(name):
    n = name.replace("void ", "").replace("gs::", "").replace("(int)", "")
    if n.startswith("render_bwd"): return "render_bwd"
    if n.startswith("render_fwd"): return "render_fwd"
    if n.startswith("preprocess_bwd"): return "preprocess_bwd"
    if n.startswith("preprocess_fwd"): return "preprocess_fwd"
    if n.startswith(("bin_scatter_kernel<1", "bin_scatter_kernel<2", "bin_count1", "bin_count2", "block_bounds")): return "tile_sort"
    if n.startswith(("bin_scatter_kernel<0", "bin_count_kernel")): return "depth_sort"
    if n.startswith("scan_"): return "scan"
    return None   # bin_scan_kernel serves both sorts (< 1 MB per launch)


acc = {}
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")]
    st = stage_of(name)
    if st is None:
        continue
    e = acc.setdefault(st, {"bytes": 0.0, "us": 0.0, "launches": 0})
    e["bytes"] += col(r, "dram__bytes_read.sum") + col(r, "dram__bytes_write.sum")
    e["launches"] += 1
    if st in ("render_bwd", "render_fwd", "preprocess_fwd", "preprocess_bwd"):
        e["pipes"] = {"issue_slot_pct": round(col(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"), 1),
                      "fma_pipe_pct": round(col(r, "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"), 1),
                      "achieved_occupancy_pct": round(col(r, "sm__warps_active.avg.pct_of_peak_sustained_active"), 1),
                      "dram_pct": round(col(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"), 1),
                      "registers": int(col(r, "launch__registers_per_thread")), "capture": capture}
res = {"source": "ncu --set full --clock-control none capture of ONE step at the bench workload (dram__bytes_read.sum + "
                 "dram__bytes_write.sum per launch, summed over the launches of a stage); see profiles/" + capture}
for st, e in acc.items():
    res[st] = {"bytes": int(e["bytes"]), "launches": e["launches"], "capture": capture}
    if "pipes" in e:
        res[st]["pipes"] = e["pipes"]
print(json.dumps(res, indent=1))
