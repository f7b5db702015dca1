#!/usr/bin/env python
"""Top stall sites of one kernel launch of an ncu report (source page, SASS view).
usage: python scripts/ncu_hot.py report.ncu-rep [launch index=0] [top=40]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; idx = int(sys.argv[2]) if len(sys.argv) > 2 else 0; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
blocks = out.split('"Kernel Name"')[1:]
rows = list(csv.reader(io.StringIO('"Kernel Name"' + blocks[idx])))
hdr = rows[1]; data = rows[2:]
ia, isrc, iall, inot = hdr.index("Address"), hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Warp Stall Sampling (Not-issued Samples)")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_")]
tot = sum(int(r[iall]) for r in data if len(r) > iall and r[iall].isdigit())
print("total samples", tot, "instructions", len(data))
base = int(data[0][ia], 16)
items = []
for r in data:
    if len(r) <= iall or not r[iall].isdigit(): continue
    st = {hdr[i][6:]: int(r[i]) for i in stall_cols if r[i].isdigit() and int(r[i]) > 0}
    items.append((int(r[iall]), int(r[ia], 16) - base, r[isrc].strip(), st))
for n, a, src, st in sorted(items, reverse=True)[:top]:
    main = sorted(st.items(), key=lambda x: -x[1])[:3]
    print("%6d %5.2f%%  %05x  %-60s %s" % (n, 100.0 * n / tot, a, src[:60], main))
# by stall reason over the whole kernel
agg = {}
for n, a, src, st in items:
    for k, v in st.items(): agg[k] = agg.get(k, 0) + v
print(sorted(agg.items(), key=lambda x: -x[1]))
