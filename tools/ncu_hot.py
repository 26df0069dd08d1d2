#!/usr/bin/env python
"""Print the hottest SASS instructions of one kernel from an ncu report's source page.
usage: tools/ncu_hot.py report.ncu-rep kernel_regex [top_n] [--all] [--name-has=SUBSTRING]
(kernel_regex matches the base function name; --name-has picks the first launch whose demangled name, template
arguments included, contains SUBSTRING)"""
import csv, subprocess, sys, io
rep, rx = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 and sys.argv[3].isdigit() else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{rx}"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
# first (matching) kernel only
has = next((a.split("=", 1)[1] for a in sys.argv if a.startswith("--name-has=")), None)
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
pick = next((n for n, i in enumerate(starts) if has is None or has in rows[i][1]), 0)
rows = rows[starts[pick]:(starts[pick + 1] if len(starts) > pick + 1 else None)]
hdr = rows[1]
ia, isrc, iex, ismp = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
body = [r for r in rows[2:] if len(r) > ismp]
tot_ex = sum(int(r[iex]) for r in body); tot_s = sum(int(r[ismp]) for r in body)
print(f"kernel {rows[0][1]}: {len(body)} SASS instr, executed warp-instr {tot_ex}, samples {tot_s}")
if "--all" in sys.argv:
    for n, r in enumerate(body):
        print(f"{n:4d} {int(r[iex]):12d} {100*int(r[iex])/tot_ex:5.2f}% smp {100*int(r[ismp])/max(tot_s,1):5.2f}%  {r[isrc].strip()[:110]}")
else:
    for n, r in sorted(enumerate(body), key=lambda x: -int(x[1][ismp]))[:top]:
        print(f"{n:4d} {int(r[iex]):12d} {100*int(r[iex])/tot_ex:5.2f}% smp {100*int(r[ismp])/max(tot_s,1):5.2f}%  {r[isrc].strip()[:110]}")
